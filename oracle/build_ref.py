"""Recipe for oracle/_ref/: the reference's own ``mfm`` package as the timed CPU arm of bench.py.

    python oracle/build_ref.py            (build container only: needs /root/reference)

The reference is pure Python, so "building" it means byte-compiling its four live source files
(Barra-master/mfm/{__init__,MFM,CrossSection,utils}.py — NOT the stale copies under build/, dist/, mfm-1.0/,
SURVEY.md §2) where they lie into code objects under oracle/_ref/mfm/ (``.pyc`` content in files named ``*.bc``: snapshots of
the repository that skip ``*.pyc`` still carry them).  Nothing is edited and
no source text is copied: the directory holds compiled code objects only, is git-ignored, and travels to the GPU
box with the snapshot like the built ``.so`` files.  oracle/ref_loader.py imports it there with the
third-party stand-ins of oracle/ref_stubs.py.  A manifest records the SHA-256 of every source compiled.
"""
import hashlib
import json
import os
import py_compile
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/Barra-master/mfm"
OUT = os.path.join(HERE, "_ref")
FILES = ["__init__.py", "MFM.py", "CrossSection.py", "utils.py"]


def build(quiet=False):
    if not os.path.isdir(SRC):
        if not quiet:
            print("reference tree not present: keeping the prebuilt oracle/_ref/ as it is")
        return None
    dst = os.path.join(OUT, "mfm")
    os.makedirs(dst, exist_ok=True)
    manifest = {"python": sys.version.split()[0], "files": {}}
    for f in FILES:
        src = os.path.join(SRC, f)
        py_compile.compile(src, cfile=os.path.join(dst, f[:-3] + ".bc"), dfile="reference:Barra-master/mfm/" + f, doraise=True)
        manifest["files"][f] = hashlib.sha256(open(src, "rb").read()).hexdigest()
    with open(os.path.join(OUT, "MANIFEST.json"), "w") as fh:
        json.dump(manifest, fh, indent=1)
    return OUT


if __name__ == "__main__":
    print(build())
