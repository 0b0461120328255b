#!/usr/bin/env python
"""(Re)install the unmodified reference under ``baseline/_ref`` -- the offline install recorded in DESIGN.md
section 5.  ``baseline/_ref`` is git-ignored, so a fresh checkout (or a replaced machine) lacks it;
``__graft_entry__.build()`` calls ``ensure()`` so that ``bench.py --impl reference`` keeps working.

  pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref <copy of /root/reference>
  + the ``models`` / ``data_utils`` sub-packages the reference's setup.py leaves out of its wheel.
No reference file is edited."""
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
TARGET = os.path.join(HERE, "_ref")
MARKER = os.path.join(TARGET, "CommEfficient", "fed_aggregator.py")


def installed() -> bool:
    return os.path.exists(MARKER) and os.path.isdir(os.path.join(TARGET, "CommEfficient", "models"))


def ensure(verbose: bool = False) -> bool:
    """True when ``baseline/_ref`` is usable afterwards.  Does nothing if it already is, or if the
    reference sources are not on this machine (e.g. on the GPU box, which receives ``_ref`` as files)."""
    if installed():
        return True
    if not os.path.isdir(os.path.join(REF_SRC, "CommEfficient")):
        return False
    tmp = tempfile.mkdtemp(prefix="ref_src_")
    try:
        src = os.path.join(tmp, "reference")
        shutil.copytree(REF_SRC, src)          # the source tree is read-only; pip builds in a copy
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links",
               "/opt/wheelhouse", "--target", TARGET, src]
        res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if res.returncode != 0:
            res = subprocess.run(cmd[:-1] + ["--no-deps", src], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if verbose or res.returncode != 0:
            print(res.stdout[-2000:])
        if res.returncode != 0:
            return False
        for sub in ("models", "data_utils"):
            dst = os.path.join(TARGET, "CommEfficient", sub)
            if not os.path.isdir(dst):
                shutil.copytree(os.path.join(REF_SRC, "CommEfficient", sub), dst)
        return installed()
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    ok = ensure(verbose=True)
    print("baseline/_ref:", "installed" if ok else "NOT available")
    sys.exit(0 if ok else 1)
