"""Install the UNMODIFIED reference (izmttk/nano-vllm-x) into baseline/_ref (git-ignored, travels to the GPU box).

Step 1 is the contract's offline pip install.  The reference's pyproject.toml lists
``packages = ["nanovllmx"]`` only (pyproject.toml:27-29), so the wheel carries the top-level
package but none of its sub-packages (core/, layers/, models/, distributed/, entrypoints/): an
upstream packaging bug.  Step 2 completes the installed tree with the missing sub-package
directories, byte for byte from /root/reference — nothing is edited.  The import fix the
reference's own bench.py needs (``from core.llm`` -> ``nanovllmx.llm``, bench.py:4-5) and the
tokenizer stub live in benchmarks/ref_engine.py, outside this tree.

Run in the build container only (/root/reference does not exist on the GPU box).
"""
import filecmp
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference"
DST = os.path.join(ROOT, "baseline", "_ref")


def install(verbose: bool = True) -> str:
    if not os.path.isdir(REF):
        raise RuntimeError(f"{REF} is absent: baseline/_ref can only be (re)built in the build container")
    os.makedirs(DST, exist_ok=True)
    src = "/tmp/_nanovllmx_ref_src"
    shutil.rmtree(src, ignore_errors=True)
    shutil.copytree(REF, src, ignore=shutil.ignore_patterns(".git"))      # the build writes egg-info next to the sources
    cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--find-links", "/opt/wheelhouse",
           "--no-deps", "--upgrade", "--target", DST, src]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("pip install of the reference failed:\n" + r.stdout[-2000:] + r.stderr[-2000:])
    pkg_src, pkg_dst = os.path.join(REF, "nanovllmx"), os.path.join(DST, "nanovllmx")
    added = []
    for name in sorted(os.listdir(pkg_src)):
        s, d = os.path.join(pkg_src, name), os.path.join(pkg_dst, name)
        if os.path.isdir(s) and name != "__pycache__" and not os.path.isdir(d):
            shutil.copytree(s, d, ignore=shutil.ignore_patterns("__pycache__"))
            added.append(name)
    cmp = filecmp.dircmp(pkg_src, pkg_dst, ignore=["__pycache__"])
    assert not cmp.left_only and not cmp.diff_files, (cmp.left_only, cmp.diff_files)
    if verbose:
        print(f"reference installed into {DST}; sub-packages missing from the wheel and completed from source: {added}")
    return DST


if __name__ == "__main__":
    install()
