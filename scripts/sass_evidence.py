"""Per-kernel SASS mnemonic counts that prove the tcgen05 / TMEM / TMA paths (run on CPU: nvcc + cuobjdump).
Writes profiles/sass_tcgen05_evidence.txt."""
import collections, os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = ["gemm.cu", "gemm_prefill.cu", "attn_prefill_tc.cu", "attention.cu", "tp_nvls.cu", "p2p.cu",
       "elementwise.cu", "sampling.cu"]
WANT = ["UTCHMMA", "UTCQMMA", "UTCCP", "UBLKCP", "UTMALDG", "LDTM", "STTM", "UTCBAR", "SYNCS", "USETMAXREG", "UTMAPF", "UBLKPF",
        "LDGMC", "STGMC"]
EXCERPT = ["UTCHMMA", "UTCQMMA", "UTCCP", "UBLKCP", "UTMALDG", "LDTM", "STTM", "UTCBAR", "LDGMC", "STGMC", "UBLKPF"]
out = ["# cuobjdump -sass of the sm_100a objects: count of tensor-core / TMEM / TMA mnemonics per kernel",
       "# UTCHMMA/UTCQMMA = tcgen05.mma (f16 / f8f6f4), LDTM/STTM = tcgen05.ld/st, UBLKCP = cp.async.bulk,",
       "# UTMALDG = cp.async.bulk.tensor (TMA tensor map), UBLKPF = cp.async.bulk.prefetch.L2, SYNCS = mbarrier ops"]
shown, excerpts = {}, collections.OrderedDict()
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
                cur = m.group(1); counts[cur] = collections.Counter(); shown[cur] = collections.Counter(); continue
            if cur:
                for w in WANT:
                    if re.search(r"\b" + w + r"[\w.]*", line): counts[cur][w] += 1
                for w in EXCERPT:      # the first two instructions of each kind, with their operands
                    if re.search(r"\b" + w + r"[\w.]*", line) and shown[cur][w] < 2:
                        shown[cur][w] += 1
                        ins = re.sub(r"/\*[0-9a-fx]+\*/", "", line).strip().rstrip(";").strip()
                        excerpts.setdefault((f, cur), []).append(ins)
        for k, c in counts.items():
            name = subprocess.check_output(["c++filt", k], text=True).strip()
            name = re.sub(r"\(.*", "", name)
            out.append(f"{f}: {name}  " + " ".join(f"{w}={c[w]}" for w in WANT if c[w]))
open(os.path.join(ROOT, "profiles", "sass_tcgen05_evidence.txt"), "w").write("\n".join(out) + "\n")
print("\n".join(out))

lines = ["# SASS excerpts (cuobjdump -sass of the sm_100a objects): for every kernel the first two instructions of each",
         "# tensor-core / tensor-memory / TMA / multimem kind, with operands.  UTCHMMA / UTCQMMA = tcgen05.mma kind::f16 /",
         "# kind::f8f6f4 and kind::mxf8f6f4.block_scale (the latter carries two extra tmem[] scale operands), UTCCP =",
         "# tcgen05.cp (scale factors -> tensor memory), LDTM / STTM = tcgen05.ld / st, UBLKCP = cp.async.bulk,",
         "# UTMALDG = cp.async.bulk.tensor, UBLKPF = cp.async.bulk.prefetch.L2, LDGMC / STGMC = multimem.ld_reduce / multimem.st."]
for (f, k), ins in excerpts.items():
    name = re.sub(r"\(.*", "", subprocess.check_output(["c++filt", k], text=True).strip())
    lines.append("")
    lines.append(f"## {f}: {name}")
    lines += ["    " + i for i in ins]
open(os.path.join(ROOT, "profiles", "sass_excerpts.txt"), "w").write("\n".join(lines) + "\n")
