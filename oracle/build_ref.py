"""Recipe for `oracle/_ref/`: the UNMODIFIED reference package, staged so that it can travel to the GPU box.

TEST / MEASUREMENT INFRASTRUCTURE, NOT PRODUCT CODE.  The reference (torchmetrics 1.3.0dev) is pure Python: there
is nothing to compile, "building" it means putting its package directory where Python can import it.  This script
packs

    /root/reference/src/torchmetrics/          ->  oracle/_ref/reference.zip : torchmetrics/          (byte for byte)
    oracle/shim/lightning_utilities/           ->  oracle/_ref/reference.zip : lightning_utilities/   (our stand-in
                                                   for the un-vendored dependency lightning-utilities
                                                   >=0.8.0,<0.10.0, requirements/base.txt:9)

into ONE archive that Python imports from directly (zipimport: the archive goes on `sys.path`), and writes
`oracle/_ref/MANIFEST.json` with a sha256 over the packed reference files, so that a reader can check that nothing
was edited on the way.  A second archive, `oracle/_ref/reference_tests.zip`, holds the reference's own unit tests of
this path (`tests/unittests/{helpers,classification,bases}`), which `oracle/acceptance.py` runs against the drop-in
package on the GPU box.  `oracle/_ref/` is git-ignored (no reference source enters the history) but not
gpurun-ignored: like the built `.so` it ships with the working tree.  `/root/reference` does not exist on the GPU
box; there `bench.py --impl reference`, the `cpu_baseline` leg and the reference-on-CUDA column import the staged
copy through `oracle/ref_loader.py`.  Run by `__graft_entry__.build()` whenever `/root/reference` is present.
"""
import hashlib
import json
import shutil
import sys
import zipfile
from pathlib import Path

HERE = Path(__file__).resolve().parent
SRC = Path("/root/reference/src/torchmetrics")
SHIM = HERE / "shim" / "lightning_utilities"
DEST = HERE / "_ref"
ARCHIVE = DEST / "reference.zip"
TESTS_SRC = Path("/root/reference/tests/unittests")
TESTS_ARCHIVE = DEST / "reference_tests.zip"  # the reference's own unit tests of the hot path (oracle/acceptance.py)
TEST_DIRS = ("", "helpers", "classification", "bases")


def tree_digest(root: Path) -> str:
    h = hashlib.sha256()
    for p in sorted(root.rglob("*.py")):
        h.update(str(p.relative_to(root)).encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def build(force: bool = False) -> Path:
    if not SRC.exists():
        raise RuntimeError(f"{SRC} is not present: oracle/_ref can only be staged where the reference is mounted")
    want = tree_digest(SRC)
    manifest = DEST / "MANIFEST.json"
    if not force and manifest.exists() and ARCHIVE.exists() and TESTS_ARCHIVE.exists():
        try:
            if json.loads(manifest.read_text()).get("sha256_reference_py") == want:
                return ARCHIVE
        except Exception:  # noqa: BLE001
            pass
    if DEST.exists():
        shutil.rmtree(DEST)
    DEST.mkdir(parents=True)
    h, files = hashlib.sha256(), 0
    with zipfile.ZipFile(ARCHIVE, "w", zipfile.ZIP_DEFLATED) as z:
        for p in sorted(SRC.rglob("*")):
            if p.is_file() and "__pycache__" not in p.parts and p.suffix != ".pyc":
                z.write(p, Path("torchmetrics") / p.relative_to(SRC))
                if p.suffix == ".py":
                    h.update(str(p.relative_to(SRC)).encode())
                    h.update(p.read_bytes())
                    files += 1
        for p in sorted(SHIM.rglob("*.py")):
            z.write(p, Path("lightning_utilities") / p.relative_to(SHIM))
    assert h.hexdigest() == want, "packed copy differs from the reference"
    n_tests = 0
    with zipfile.ZipFile(TESTS_ARCHIVE, "w", zipfile.ZIP_DEFLATED) as z:
        for d in TEST_DIRS:
            for p in sorted((TESTS_SRC / d).glob("*.py")):
                z.write(p, Path("unittests") / p.relative_to(TESTS_SRC))
                n_tests += 1
    manifest.write_text(json.dumps({
        "source": str(SRC), "sha256_reference_py": want, "files": files, "archive": ARCHIVE.name,
        "tests_archive": TESTS_ARCHIVE.name, "test_files": n_tests,
        "shim": "oracle/shim/lightning_utilities (stand-in for lightning-utilities >=0.8.0,<0.10.0)",
    }, indent=1))
    return ARCHIVE


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
