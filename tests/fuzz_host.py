"""Mutation fuzzing of the host-side entry points that take bytes from outside the process: ksb_index_from_blob (+ the shard
builder and the row cuts on what it accepts), ksb_index_load on damaged index files, ksb_table_read_tsv / _sidecar on damaged
tables.  Nothing is asserted about WHAT is refused: the point is that no input crashes or trips a sanitizer, so run it
against the sanitized build -- `FUZZ=1 tests/sanitize_cpu.sh` -- or plainly: python tests/fuzz_host.py [seed].
(Not collected by pytest; tests/test_host_side.py::test_from_blob_scans_the_member_values pins the cases it found.)"""
import sys, os, numpy as np, struct, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from kspider_b200 import synth, Index
from kspider_b200.capi import KsbError, lib
import ctypes as C
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)
L = lib()
# --- 1. blobs: mutate descriptor words and payload words, the validator has to refuse or the accessors have to survive
def exercise(ix):
    ix.info
    for w in (1, 2, 3):
        try:
            sh = ix.shard(w - 1, w, 2); sh.info; del sh
        except KsbError: pass
    cuts = (C.c_uint64 * 5)()
    L.ksb_index_row_cuts(ix._h, 4, cuts)
cases = []
s1 = synth.fracminhash_index(n_clades=6, clade_size=40, sketch=300, core=200, seed=3)
cases.append(Index.from_csr(s1.offsets, s1.members, s1.weights, s1.total_ids, s1.total_vals, host_threads=2))
s2 = synth.zipf_index(n_samples=300, n_colors=4000, s_max=200, seed=5)
cases.append(Index.from_csr(s2.offsets, s2.members, s2.weights, s2.total_ids, s2.total_vals, host_threads=2))
os.environ["KSPIDER_GRAM_FORM"] = "csr"
cases.append(Index.from_csr(s1.offsets, s1.members, s1.weights, s1.total_ids, s1.total_vals, host_threads=2))
del os.environ["KSPIDER_GRAM_FORM"]
n_ok = n_bad = 0
for ix in cases:
    d0 = np.frombuffer(ix.descriptor(), np.uint8).copy(); p0 = ix.payload().copy()
    print("desc", d0.size, "payload", p0.size, ix.info.gram_form)
    for it in range(int(os.environ.get('FUZZ_BLOBS', 1500))):
        d = d0.copy(); p = p0.copy()
        kind = rng.integers(0, 4)
        if kind == 0:      # a descriptor word gets a random / boundary value
            w = rng.integers(1, d.size // 8)
            v = [0, 1, 0xFFFFFFFF, 2**63, 2**64 - 1, int(rng.integers(0, 2**40)), (int(d.view(np.uint64)[w]) + int(rng.integers(-3, 4))) & (2**64-1)][int(rng.integers(0, 7))]
            d.view(np.uint64)[w] = np.uint64(v)
        elif kind == 1:    # a payload u32 word
            w = rng.integers(0, p.size // 4)
            p.view(np.uint32)[w] = np.uint32(rng.choice([0, 0xFFFFFFFF, int(rng.integers(0, 2**32)), (int(p.view(np.uint32)[w]) + 1) & 0xFFFFFFFF]))
        elif kind == 2:    # a run of payload bytes randomised
            a = rng.integers(0, p.size - 64); p[a:a + 64] = rng.integers(0, 256, 64, dtype=np.uint8)
        else:              # truncated payload
            p = p[: rng.integers(0, p.size)].copy()
        try:
            cp = Index.from_blob(d.tobytes(), p)
            exercise(cp); n_ok += 1
            del cp
        except KsbError:
            n_bad += 1
print("blobs accepted", n_ok, "refused", n_bad)
# --- 2. index files: byte flips and truncations
td = tempfile.mkdtemp()
prefix = os.path.join(td, "ix")
synth.write_index_files(s1, prefix)
files = {suf: open(prefix + suf, "rb").read() for suf in ("_color_to_sources.bin", "_color_count.bin", "_groupID_to_kmerCount.bin")}
n_ok = n_bad = 0
for it in range(1500):
    suf = list(files)[rng.integers(0, 3)]
    b = bytearray(files[suf])
    kind = rng.integers(0, 3)
    if kind == 0:
        for _ in range(rng.integers(1, 4)): b[rng.integers(0, len(b))] = rng.integers(0, 256)
    elif kind == 1:
        b = b[: rng.integers(0, len(b))]
    else:
        a = rng.integers(0, len(b) - 8); b[a:a+8] = struct.pack("<Q", [0, 2**32, 2**64-1, 2**31, int(rng.integers(0, 2**20))][int(rng.integers(0, 5))])
    for s_, data in files.items(): open(prefix + s_, "wb").write(bytes(b) if s_ == suf else data)
    try:
        ix = Index.load(prefix, 3); exercise(ix); n_ok += 1; del ix
    except KsbError:
        n_bad += 1
print("files accepted", n_ok, "refused", n_bad)
# --- 3. TSV / side-car readers
from kspider_b200.api import RECORD_DTYPE
rec = np.zeros(3000, RECORD_DTYPE)
rec["source_1"] = rng.integers(1, 200, 3000); rec["source_2"] = rng.integers(1, 200, 3000); rec["shared_kmers"] = 5
for c in ("min_containment", "avg_containment", "max_containment"): rec[c] = rng.random(3000, dtype=np.float32)
tsv = os.path.join(td, "t.tsv"); side = os.path.join(td, "t.ksb")
assert L.ksb_write_tsv_records(rec.ctypes.data, 3000, tsv.encode(), 3, 0) == 0
assert L.ksb_write_sidecar_records(rec.ctypes.data, 3000, side.encode(), 0, 0, 0) == 0
t0 = open(tsv, "rb").read(); k0 = open(side, "rb").read()
n_ok = n_bad = 0
for it in range(1500):
    b = bytearray(t0); k = bytearray(k0)
    kind = rng.integers(0, 4)
    if kind == 0:
        for _ in range(rng.integers(1, 6)): b[rng.integers(0, len(b))] = rng.choice([9, 10, 32, 43, 45, 46, 48, 57, 101, 0, 255, 13])
    elif kind == 1: b = b[: rng.integers(0, len(b))]
    elif kind == 2:
        for _ in range(rng.integers(1, 4)): k[rng.integers(0, len(k))] = rng.integers(0, 256)
    else: k = k[: rng.integers(0, len(k))]
    open(tsv, "wb").write(bytes(b)); open(side, "wb").write(bytes(k))
    for col, col2 in ((5, 0), (3, 5)):
        for reader in (0, 1):
            t = C.c_void_p()
            rc = L.ksb_table_read_tsv(tsv.encode(), col, col2, None, 3, C.byref(t)) if reader == 0 else L.ksb_table_read_sidecar(side.encode(), col, col2, 3, C.byref(t))
            if rc == 0: n_ok += 1; L.ksb_table_free(t)
            else: n_bad += 1
    L.ksb_sidecar_matches(side.encode(), tsv.encode())
print("tables accepted", n_ok, "refused", n_bad)
# --- 4. the colored_kDataFrame file set (ksb_index_load_colored parses all five files on the host before any device work)
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CK = [".phmap", "colors.intvectors", "colorTableConfig.json", ".namesMap", ".extra"]
ck0 = {suf: open(os.path.join(GOLD, "ckf_small" + suf), "rb").read() for suf in CK}
cprefix = os.path.join(td, "ck")
codes = {}
for it in range(1500):
    suf = CK[int(rng.integers(0, len(CK)))]
    b = bytearray(ck0[suf])
    kind = int(rng.integers(0, 4))
    if kind == 0 and len(b):
        for _ in range(int(rng.integers(1, 5))): b[int(rng.integers(0, len(b)))] = int(rng.integers(0, 256))
    elif kind == 1:
        b = b[: int(rng.integers(0, len(b) + 1))]
    elif kind == 2 and len(b) > 8:
        a = int(rng.integers(0, len(b) - 8)); b[a:a + 8] = struct.pack("<Q", [0, 2**32, 2**64 - 1, 2**31, int(rng.integers(0, 2**20))][int(rng.integers(0, 5))])
    elif len(b):   # a digit / separator of the text files changed
        for _ in range(int(rng.integers(1, 4))): b[int(rng.integers(0, len(b)))] = [32, 10, 48, 57, 45, 9, 0][int(rng.integers(0, 7))]
    for s_, data in ck0.items(): open(cprefix + s_, "wb").write(bytes(b) if s_ == suf else data)
    h = C.c_void_p()
    rc = L.ksb_index_load_colored(cprefix.encode(), 2, -1, C.byref(h))
    codes[rc] = codes.get(rc, 0) + 1
    if rc == 0: L.ksb_index_free(h)
print("colored file sets by return code", dict(sorted(codes.items())))
# --- 5. ksb_index_from_csr on edge-shaped inputs (empty and one-member colours, zero weights, ids up to 2^32 - 1, a duplicate
# member, offsets out of order); whatever it accepts must survive its own validator as a blob and as shards of 1, 2 and 5
os.environ.pop("KSPIDER_GRAM", None); os.environ.pop("KSPIDER_GRAM_FORM", None)
ok = bad = 0
for it in range(int(os.environ.get('FUZZ_CSR', 150))):
    nc = int(rng.integers(0, 40))
    mode = int(rng.integers(0, 5))
    hi = [8, 300, 70000, 2**20, 2**32][mode]
    sizes = rng.integers(0, [6, 70, 200][int(rng.integers(0, 3))], nc)
    if mode >= 3: sizes = np.minimum(sizes, 5)
    cols = []
    for s in sizes:
        s = int(min(s, hi))
        m = rng.choice(hi, size=s, replace=False) if hi <= 70000 else np.unique(rng.integers(0, hi, s, dtype=np.uint64))
        if rng.random() < 0.02 and s > 1: m[0] = m[1]   # a duplicate: must be refused
        cols.append(np.asarray(m, dtype=np.uint32))
    offsets = np.concatenate([[0], np.cumsum([c.size for c in cols])]).astype(np.uint64)
    members = np.concatenate(cols).astype(np.uint32) if cols else np.zeros(0, np.uint32)
    weights = rng.integers(0, [2, 200, 2**32][int(rng.integers(0, 3))], nc, dtype=np.uint64).astype(np.uint32)
    nt = int(rng.integers(0, 30))
    tid = rng.integers(0, hi, nt, dtype=np.uint64).astype(np.uint32)
    tv = rng.integers(0, 2**32, nt, dtype=np.uint64).astype(np.uint32)
    if rng.random() < 0.03 and nc > 2: offsets[1], offsets[2] = offsets[2], offsets[1]
    for env in ({}, {"KSPIDER_GRAM": "16"}, {"KSPIDER_GRAM": "16", "KSPIDER_GRAM_FORM": "csr"}):
        os.environ.pop("KSPIDER_GRAM", None); os.environ.pop("KSPIDER_GRAM_FORM", None); os.environ.update(env)
        try:
            ix = Index.from_csr(offsets=offsets, members=members, weights=weights, total_ids=tid, total_vals=tv, host_threads=int(rng.integers(1, 5)))
            i = ix.info
            cp = Index.from_blob(ix.descriptor(), ix.payload())
            for w in (1, 2, 5):
                for r in range(w):
                    sh = ix.shard(r, w, 2); Index.from_blob(sh.descriptor(), sh.payload())
            ok += 1
        except KsbError as e:
            bad += 1
            if "blob" in str(e): print("BLOB REFUSED", e, env, nc, mode); raise
print("from_csr inputs accepted", ok, "refused", bad)
