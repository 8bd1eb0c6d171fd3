"""Per-kernel SASS mnemonic counts that prove the tcgen05 / TMEM / TMA paths (run on CPU: nvcc + cuobjdump).
Writes profiles/sass_tcgen05_evidence.txt."""
import collections, os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = ["gemm.cu", "gemm_prefill.cu", "attn_prefill_tc.cu", "attention.cu", "p2p.cu", "elementwise.cu", "sampling.cu"]
WANT = ["UTCHMMA", "UTCQMMA", "UBLKCP", "UTMALDG", "LDTM", "STTM", "UTCBAR", "SYNCS", "USETMAXREG", "UTMAPF", "UBLKPF"]
out = ["# cuobjdump -sass of the sm_100a objects: count of tensor-core / TMEM / TMA mnemonics per kernel",
       "# UTCHMMA/UTCQMMA = tcgen05.mma (f16 / f8f6f4), LDTM/STTM = tcgen05.ld/st, UBLKCP = cp.async.bulk,",
       "# UTMALDG = cp.async.bulk.tensor (TMA tensor map), UBLKPF = cp.async.bulk.prefetch.L2, SYNCS = mbarrier ops"]
with tempfile.TemporaryDirectory() as tmp:
    for f in SRC:
        obj = os.path.join(tmp, f + ".o")
        subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-I" + os.path.join(ROOT, "include"),
                               "-I" + os.path.join(ROOT, "csrc"), "-c", os.path.join(ROOT, "csrc", f), "-o", obj])
        sass = subprocess.check_output(["cuobjdump", "-sass", obj], text=True)
        cur, counts = None, collections.OrderedDict()
        for line in sass.splitlines():
            m = re.search(r"Function : (\S+)", line)
            if m:
                cur = m.group(1); counts[cur] = collections.Counter(); continue
            if cur:
                for w in WANT:
                    if re.search(r"\b" + w + r"[\w.]*", line): counts[cur][w] += 1
        for k, c in counts.items():
            name = subprocess.check_output(["c++filt", k], text=True).strip()
            name = re.sub(r"\(.*", "", name)
            out.append(f"{f}: {name}  " + " ".join(f"{w}={c[w]}" for w in WANT if c[w]))
open(os.path.join(ROOT, "profiles", "sass_tcgen05_evidence.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(out))
