"""oracle/_ref -- the reference's own tick path, compiled, so that it can run on the GPU box.

TEST / MEASUREMENT INFRASTRUCTURE.  The reference is pure Python; `/root/reference` exists only in the build container.
This recipe byte-compiles (py_compile, CPython's own compiler -- the Python counterpart of building a C reference with
gcc) exactly the modules `nsim.Simulator` needs for headless load / reset / tick -- found by importing the reference and
listing sys.modules, 43 modules -- into git-ignored `oracle/_ref/nclone/**.refc` (a .pyc image under another extension:
snapshot tools skip *.pyc; oracle/ref_baseline.py imports them with a small meta-path finder).  No reference source
is copied: only compiler output lands under oracle/_ref/, which travels to the GPU box with the snapshot like the
built .so files.  Used only by bench.py's `--impl reference` arm and `cpu_baseline` leg (oracle/ref_baseline.py), never
by the product.

    python -m oracle.make_ref          (also called from __graft_entry__.build() when /root/reference is present)
"""
import os
import py_compile
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("NCLONE_REF", "/root/reference")
OUT = os.path.join(HERE, "_ref")

_PROBE = r"""
import os, sys
os.environ["HOME"] = "/tmp/nclone_ref_home"; os.makedirs(os.environ["HOME"], exist_ok=True)
sys.dont_write_bytecode = True
sys.path.insert(0, sys.argv[1])
from nclone.nsim import Simulator
from nclone.sim_config import SimConfig
s = Simulator(SimConfig(enable_anim=False, log_data=False, debug=False))
md = [0] * 1245; md[1231] = 22; md[1232] = 22; md[1233] = 1; md[1156] = 1
md[1235:1240] = [3, 40, 22, 0, 0]; md[1240:1245] = [4, 30, 22, 0, 0]
s.load(md); s.tick(1, 0); s.reset(); s.fast_reset(); s.tick(0, 1)
for name, m in sorted(sys.modules.items()):
    f = getattr(m, "__file__", None)
    if name.split(".")[0] == "nclone" and f and f.endswith(".py"):
        print(name, f)
"""


def available():
    return os.path.isfile(os.path.join(OUT, "nclone", "nsim.refc"))


def build(force=False):
    """-> number of modules compiled (0 when the reference tree is absent, e.g. on the GPU box)."""
    if not os.path.isdir(REF_ROOT):
        return 0
    if available() and not force:
        return -1
    r = subprocess.run([sys.executable, "-c", _PROBE, REF_ROOT], capture_output=True, text=True, check=True)
    mods = [ln.split(" ", 1) for ln in r.stdout.strip().splitlines()]
    shutil.rmtree(OUT, ignore_errors=True)
    for name, src in mods:
        rel = os.path.relpath(src, REF_ROOT)
        dst = os.path.join(OUT, rel[:-3] + ".refc")
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        py_compile.compile(src, cfile=dst, dfile=rel, doraise=True, optimize=0)
    with open(os.path.join(OUT, "MANIFEST.txt"), "w") as f:
        f.write(f"compiled by oracle/make_ref.py with CPython {sys.version.split()[0]} from {REF_ROOT}\n")
        f.writelines(f"{name}\n" for name, _ in mods)
    return len(mods)


if __name__ == "__main__":
    n = build(force=True)
    print(f"oracle/_ref: {n} modules compiled" if n else "reference tree not present: nothing built")
