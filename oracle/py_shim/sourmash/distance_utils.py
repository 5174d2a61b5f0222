"""TEST INFRASTRUCTURE ONLY — the one sourmash call the reference makes (ks_pairwise.py:50,78-79):
containment_to_distance(containment, ksize, scaled, n_unique_kmers=...).ani

sourmash is third-party, listed nowhere with a version (not in setup.py:203-210 / requirements.txt), absent
from /root/reference and from this image, so its arithmetic is restated from the published algorithm of
sourmash >= 4.4 (sourmash/distance_utils.py: containment_to_distance with the defaults confidence=0.95,
estimate_ci=False, prob_threshold=1e-3; ANIResult.__post_init__ / .ani): "parity unpinned" for THIS formula.
What running the reference through this shim pins is everything ks_pairwise.py itself does: which text
columns it reads, float() of the printed values, the mean, the f-string repr, header and row order.
"""
from dataclasses import dataclass


@dataclass
class ANIResult:
    dist: float
    p_nothing_in_common: float = 0.0
    p_threshold: float = 1e-3
    size_is_inaccurate: bool = False

    def __post_init__(self):
        if not (0 <= self.dist <= 1 and 0 <= self.p_threshold <= 1):
            raise ValueError(f"Error: distance value {self.dist :.4f} and p_threshold value "
                             f"{self.p_threshold :.4f} must both be between 0 and 1.")

    @property
    def ani(self):
        if self.size_is_inaccurate:
            return None
        return 1 - self.dist


def containment_to_distance(containment, ksize, scaled, n_unique_kmers=None, sequence_len_bp=None,
                            confidence=0.95, estimate_ci=False, prob_threshold=1e-3):
    if containment == 0:
        point_estimate = 1.0
    elif containment == 1:
        point_estimate = 0.0
    else:
        point_estimate = 1.0 - containment ** (1.0 / ksize)
    return ANIResult(point_estimate, 0.0, prob_threshold)
