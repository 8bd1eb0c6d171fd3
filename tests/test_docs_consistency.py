"""The documents name flags, environment switches and files; they must exist (a renamed flag or a deleted profile must
not survive in the prose)."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DOCS = ["README.md", "DESIGN.md", "INTEGRATION.md", "BASELINE.md", "docs/DEPLOY.md", "docs/LOAD_BALANCING.md",
        "docs/FAULT_TOLERANCE.md", "docs/TROUBLESHOOTING.md", "docs/BENCHMARKS.md", "docs/ARCHITECTURE.md", "profiles/README.md"]


def _read(rel):
    with open(os.path.join(ROOT, rel)) as f:
        return f.read()


def _sources():
    out = []
    for base in ("src", "csrc", "scripts", "global_capstone_design_distributed-inference-of-llms-over-the-internet_b200",
                 "baseline"):
        for d, _, files in os.walk(os.path.join(ROOT, base)):
            if "_ref" in d or "__pycache__" in d:
                continue
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                    out.append(open(os.path.join(d, f), errors="replace").read())
    out.append(_read("bench.py"))
    return "\n".join(out)


def test_cli_flags_named_in_the_docs_exist():
    from src.main import build_parser

    known = {a for act in build_parser()._actions for a in act.option_strings}
    # flags of other programs that the docs legitimately mention
    other = {"--gpus", "--steps", "--warmup", "--impl", "--weights", "--hop", "--head-split", "--no-ttft", "--timeout",
             "--trace", "--nproc-per-node", "--master-addr", "--master-port", "--nnodes", "--set", "--clock-control",
             "--import-source", "--page", "--csv", "--kill_stage", "--kill_at_token", "--hard", "--out", "--clients",
             "--yes", "--metrics", "--standalone", "--local-addr", "--base_port", "--port_base", "--graph-profiling",
             "--sessions", "--group", "--tp-mode", "--prompt-len", "--linger_ms", "--soft", "--no-index",
             "--no-build-isolation", "--find-links", "--target", "--no-deps", "--log-file", "--launch-skip"}
    missing = {}
    for doc in DOCS:
        for flag in set(re.findall(r"(?<![\w-])--[a-z][a-z0-9_-]+", _read(doc))):
            if flag not in known and flag not in other:
                missing.setdefault(doc, []).append(flag)
    assert not missing, f"flags in the docs that no parser knows: {missing}"


def test_environment_switches_named_in_the_docs_are_read_somewhere():
    src = _sources()
    missing = {}
    for doc in DOCS:
        for var in set(re.findall(r"\b(B200_[A-Z0-9_]+|PETALS_B200_[A-Z0-9_]+)\b", _read(doc))):
            if var.endswith("_") or var in ("B200_E", "B200_OK", "B200_ELIMIT", "B200_ECUDA", "B200_EINVAL", "B200_BF16",
                                            "B200_PROFILING", "B200_MICROARCH", "B200_TRACE"):
                continue
            if var not in src:
                missing.setdefault(doc, []).append(var)
    assert not missing, f"environment switches in the docs that nothing reads: {missing}"


def test_files_named_in_the_docs_exist():
    missing = {}
    for doc in DOCS:
        text = _read(doc)
        for m in set(re.findall(r"`((?:profiles|scripts|tests|docs|csrc|src|include|baseline)/[A-Za-z0-9_./{},*-]+)`", text)):
            m = m.split("::")[0].split(":")[0].rstrip(".,")
            if any(ch in m for ch in "*{"):
                continue
            if not os.path.exists(os.path.join(ROOT, m)):
                missing.setdefault(doc, []).append(m)
        if doc == "profiles/README.md":
            for m in set(re.findall(r"`([A-Za-z0-9_.-]+\.(?:json|csv|md|txt|log))`", text)):
                if not os.path.exists(os.path.join(ROOT, "profiles", m)) and not os.path.exists(os.path.join(ROOT, m)):
                    missing.setdefault(doc, []).append("profiles/" + m)
    assert not missing, f"files named in the docs that do not exist: {missing}"
