"""TEST INFRASTRUCTURE ONLY — stand-in for the SWIG extension `_kSpider_internal`
(src/swig_interfaces/kSpider_internal.i; swig and the CMake build are unavailable in this image), so that
the UNMODIFIED reference package /root/reference/pykSpider/kSpider2 imports and its CLI runs.
pairwise(index_prefix, user_threads) runs oracle/_ref/pairwise_ref = the UNMODIFIED src/pairwise.cpp
(oracle/Makefile); the other entry points of the .i file are not on the path and raise."""
import os
import subprocess
import sys

_REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_ref", "pairwise_ref")


def pairwise(index_prefix, user_threads):
    p = subprocess.run([_REF, str(index_prefix), str(int(user_threads))], stdout=subprocess.PIPE, text=True, check=True)
    sys.stdout.write(p.stdout)


def __getattr__(name):
    def missing(*a, **k):
        raise NotImplementedError(f"_kSpider_internal.{name} is not on the pairwise path (oracle shim)")
    return missing
