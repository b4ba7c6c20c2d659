"""Places the UNMODIFIED reference under baseline/_ref (git-ignored; it travels to the GPU box).

Step 1 is the sanctioned offline install:
    python -m pip install --no-index --no-build-isolation --find-links /opt/wheelhouse \
        --target baseline/_ref /root/reference
The reference has no setup.py / pyproject.toml, so pip answers "Directory '/root/reference' is
not installable" (recorded in DESIGN.md).  Step 2, the fallback, copies the source tree verbatim —
nothing is edited; bench.py --impl reference imports the reference's own modules from there.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
SRC = os.environ.get("LDNN_REFERENCE_SRC", "/root/reference")


def install(verbose=True) -> str:
    if os.path.isdir(os.path.join(DST, "Balanced All-Reduce")):
        return "present"
    if not os.path.isdir(SRC):
        return "unavailable: reference source tree not found at " + SRC
    r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
                        "/opt/wheelhouse", "--target", DST, SRC], capture_output=True, text=True)
    note = "pip: ok"
    if r.returncode != 0 or not os.path.isdir(os.path.join(DST, "Balanced All-Reduce")):
        note = "pip: " + (r.stderr.strip().splitlines() or ["failed"])[-1]
        os.makedirs(DST, exist_ok=True)
        for name in os.listdir(SRC):
            s, d = os.path.join(SRC, name), os.path.join(DST, name)
            if os.path.isdir(s):
                shutil.copytree(s, d, dirs_exist_ok=True, ignore=shutil.ignore_patterns("Graphs", "__pycache__"))
            else:
                shutil.copy2(s, d)
        note += " | fallback: verbatim copy of the source tree"
    with open(os.path.join(DST, "INSTALL_NOTE.txt"), "w") as f:
        f.write(note + "\n")
    if verbose:
        print("[install_ref]", note)
    return note


if __name__ == "__main__":
    print(install())
