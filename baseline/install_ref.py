"""Put the UNMODIFIED reference files of the hot path under baseline/_ref/ (git-ignored, but it travels to the GPU box with the
snapshot like the built .so does), so that `bench.py --impl reference` and the `cpu_baseline` leg time the reference's OWN
modules there.  The reference is not a pip package (no setup.py / pyproject.toml; only the vendored utils/nlg_eval has one), so
"install" is a copy of the sources the path needs:
    models/*.py, train_iq.py (TrainIQ.calculate_losses), utils/train_utils.py, utils/__init__.py
Run here (where /root/reference exists):  python baseline/install_ref.py        — __graft_entry__.build() does it too.
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.environ.get("B200IQ_REFERENCE_SRC", "/root/reference")
DST = os.path.join(HERE, "_ref")
FILES = ["train_iq.py", "utils/train_utils.py", "utils/__init__.py", "LICENSE"]


def install(verbose=True):
    if not os.path.isdir(os.path.join(SRC, "models")):
        if verbose:
            print("reference tree not found at %s: nothing installed" % SRC)
        return False
    os.makedirs(DST, exist_ok=True)
    shutil.copytree(os.path.join(SRC, "models"), os.path.join(DST, "models"), dirs_exist_ok=True,
                    ignore=shutil.ignore_patterns("__pycache__"))
    for f in FILES:
        s = os.path.join(SRC, f)
        if os.path.exists(s):
            os.makedirs(os.path.dirname(os.path.join(DST, f)), exist_ok=True)
            shutil.copyfile(s, os.path.join(DST, f))
    if verbose:
        n = sum(len(fs) for _, _, fs in os.walk(DST))
        print("reference hot-path sources copied to %s (%d files)" % (DST, n))
    return True


if __name__ == "__main__":
    sys.exit(0 if install() else 1)
