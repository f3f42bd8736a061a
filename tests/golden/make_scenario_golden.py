"""Regenerates tests/golden/scenario/: runs the UNMODIFIED reference scenario generator
(/root/reference/scripts/generate_config.py, which reads scripts/example.yml and writes
relative to the working directory) in a scratch directory and copies what it wrote, next
to the YAML it was given.  Run in the build container only (the GPU box has no reference)."""
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/scripts"
OUT = os.path.join(HERE, "scenario")

with tempfile.TemporaryDirectory() as tmp:
    subprocess.run([sys.executable, os.path.join(REF, "generate_config.py")], cwd=tmp, check=True,
                   stdout=subprocess.DEVNULL)
    if os.path.isdir(OUT):
        shutil.rmtree(OUT)
    shutil.copytree(os.path.join(tmp, "config"), os.path.join(OUT, "config"))
shutil.copy(os.path.join(REF, "example.yml"), os.path.join(OUT, "example.yml"))
for root, _, files in os.walk(OUT):
    for f in files:
        print(os.path.relpath(os.path.join(root, f), OUT))
