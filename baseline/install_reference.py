"""Stage the UNMODIFIED reference under ``baseline/_ref`` (git-ignored, travels with gpurun).

``pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target baseline/_ref
/root/reference`` fails ("Neither 'setup.py' nor 'pyproject.toml' found": the reference is a workshop
site, not a package — SURVEY §0), so the install is a byte-for-byte copy of ``notebooks/code`` (the
reference's whole Python surface) verified by SHA-256.  Nothing is edited; ``baseline/run_reference.py``
drives the copied scripts through their own CLI."""
import hashlib
import os
import shutil
import sys

SRC = "/root/reference/notebooks/code"
DST = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "code")


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 20), b""):
            h.update(chunk)
    return h.hexdigest()


def install(include_checkpoints: bool = False) -> str:
    if not os.path.isdir(SRC):
        raise RuntimeError("reference tree not mounted at %s" % SRC)
    os.makedirs(DST, exist_ok=True)
    n = 0
    for root, dirs, files in os.walk(SRC):
        dirs[:] = [d for d in dirs if d not in (".ipynb_checkpoints", "__pycache__") and
                   (include_checkpoints or d != "shadow_model_ckpt")]
        rel = os.path.relpath(root, SRC)
        os.makedirs(os.path.join(DST, rel), exist_ok=True)
        for f in files:
            s, d = os.path.join(root, f), os.path.join(DST, rel, f)
            shutil.copyfile(s, d)
            assert _sha(s) == _sha(d)
            n += 1
    with open(os.path.join(os.path.dirname(DST), "INSTALL_LOG.txt"), "w") as f:
        f.write("pip install of /root/reference: not installable (no setup.py/pyproject.toml)\n")
        f.write("staged %d files from %s byte-identical (sha256 verified)\n" % (n, SRC))
    return DST


if __name__ == "__main__":
    print(install("--with-checkpoints" in sys.argv))
