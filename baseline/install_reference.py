"""Install the reference into ``baseline/_ref`` (git-ignored; travels to the GPU box).

Step 1 is the prescribed offline pip install.  The reference is a bare script (no
``setup.py`` / ``pyproject.toml``), so pip answers "Directory ... is not installable";
in that case step 2 vendors the two reference files, byte for byte, next to a manifest of
their sha256 sums so ``bench.py --impl reference`` can run the UNMODIFIED script from
``baseline/_ref``.  Outcome recorded in DESIGN.md.
"""
from __future__ import annotations

import hashlib
import json
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

HERE = Path(__file__).resolve().parent
DEST = HERE / "_ref"
SRC = Path("/root/reference")


def sha256(p: Path) -> str:
    return hashlib.sha256(p.read_bytes()).hexdigest()


def main() -> int:
    if not SRC.exists():
        print(f"{SRC} not present; nothing to install (baseline/_ref kept as is)")
        return 0
    DEST.mkdir(parents=True, exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp:
        copy = Path(tmp) / "reference"
        shutil.copytree(SRC, copy)      # /root/reference is read-only; pip may write into the tree
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation",
               "--find-links", "/opt/wheelhouse", "--target", str(DEST), str(copy)]
        proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    pip_ok = proc.returncode == 0
    print(proc.stdout.strip().splitlines()[-1] if proc.stdout.strip() else "(no pip output)")
    manifest = {"pip_install_ok": pip_ok, "pip_tail": proc.stdout.strip()[-400:], "files": {}}
    if not pip_ok:
        for f in sorted(SRC.iterdir()):
            if f.is_file():
                shutil.copy2(f, DEST / f.name)
                manifest["files"][f.name] = sha256(DEST / f.name)
                assert manifest["files"][f.name] == sha256(f)
        print(f"pip cannot install a bare script; vendored {sorted(manifest['files'])} unmodified into {DEST}")
    (DEST / "MANIFEST.json").write_text(json.dumps(manifest, indent=1))
    return 0


if __name__ == "__main__":
    sys.exit(main())
