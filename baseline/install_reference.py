"""Puts the unmodified reference sources under baseline/_ref (git-ignored; travels with gpurun).

``pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref
/root/reference`` fails because the reference has no setup.py / pyproject.toml (also with
``--no-deps`` from a /tmp copy), so the two pure-Python packages it consists of (``src`` and
``petals``) are copied verbatim instead.  Nothing is edited.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
SRC = os.environ.get("REFERENCE_ROOT", "/root/reference")


def main() -> int:
    if not os.path.isdir(SRC):
        print(f"{SRC} not found; nothing to install")
        return 1
    r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
                        "/opt/wheelhouse", "--target", DST, SRC], capture_output=True, text=True)
    print("pip install:", "ok" if r.returncode == 0 else (r.stderr.strip().splitlines() or ["failed"])[-1])
    os.makedirs(DST, exist_ok=True)
    for pkg in ("src", "petals"):
        d = os.path.join(DST, pkg)
        if os.path.isdir(d):
            shutil.rmtree(d)
        shutil.copytree(os.path.join(SRC, pkg), d)
    print("copied src/ and petals/ to", DST)
    return 0


if __name__ == "__main__":
    sys.exit(main())
