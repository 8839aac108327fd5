"""Pack the UNMODIFIED reference (MarvinLvn/CPC2, /root/reference) into oracle/_ref/ -- TEST / BASELINE INFRASTRUCTURE.

    python oracle/make_ref.py            (also run by __graft_entry__.build() when /root/reference is present)

/root/reference does not exist on the GPU box, so the reference arm of bench.py (`--impl reference`, `stock_gpu`)
needs the reference to travel with the snapshot.  This script stores every `cpc/**/*.py` of the reference, byte for
byte, in ONE archive `oracle/_ref/cpc_reference.zip` plus a manifest of per-file SHA-256 sums
(`oracle/_ref/MANIFEST.json`); `oracle/ref_harness.py` puts the archive on `sys.path` (zipimport) when
/root/reference is absent.  `oracle/_ref/` is git-ignored (the history holds no reference source) but not
gpurun-ignored, so the archive ships to the GPU box like a built `.so`.  Nothing under cpc2_b200/ reads it.
"""
import hashlib
import json
import os
import sys
import zipfile

HERE = os.path.dirname(os.path.abspath(__file__))
OUT_DIR = os.path.join(HERE, "_ref")
ARCHIVE = os.path.join(OUT_DIR, "cpc_reference.zip")
MANIFEST = os.path.join(OUT_DIR, "MANIFEST.json")
REFERENCE_ROOT = os.environ.get("CPC2_REFERENCE_ROOT", "/root/reference")


def make(force=False):
    """Returns the archive path, or None when there is neither a reference tree nor an archive."""
    src = os.path.join(REFERENCE_ROOT, "cpc")
    if not os.path.isdir(src):
        return ARCHIVE if os.path.exists(ARCHIVE) else None
    files = []
    for d, _, fs in os.walk(src):
        for f in sorted(fs):
            if f.endswith(".py"):
                files.append(os.path.join(d, f))
    files.sort()
    sums = {}
    for p in files:
        with open(p, "rb") as fh:
            sums[os.path.relpath(p, REFERENCE_ROOT)] = hashlib.sha256(fh.read()).hexdigest()
    if not force and os.path.exists(ARCHIVE) and os.path.exists(MANIFEST):
        with open(MANIFEST) as fh:
            if json.load(fh).get("sha256") == sums:
                return ARCHIVE
    os.makedirs(OUT_DIR, exist_ok=True)
    with zipfile.ZipFile(ARCHIVE, "w", zipfile.ZIP_DEFLATED) as z:
        for p in files:
            z.write(p, os.path.relpath(p, REFERENCE_ROOT))
    with open(MANIFEST, "w") as fh:
        json.dump({"source": REFERENCE_ROOT, "files": len(files), "sha256": sums}, fh, indent=1, sort_keys=True)
    return ARCHIVE


if __name__ == "__main__":
    path = make(force="--force" in sys.argv)
    print(path if path else "no reference tree at " + REFERENCE_ROOT)
