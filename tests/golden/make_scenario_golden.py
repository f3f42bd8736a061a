"""Regenerates tests/golden/scenario/config: runs the UNMODIFIED reference scenario generator
(/root/reference/scripts/generate_config.py; it opens `<its own directory>/<--config-file>` and
writes relative to the working directory) on tests/golden/scenario/crowd.yml in a scratch
directory and copies what it wrote.  Run in the build container only (the GPU box has no
reference tree)."""
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/scripts"
OUT = os.path.join(HERE, "scenario")

with tempfile.TemporaryDirectory() as tmp:
    rel = os.path.relpath(os.path.join(OUT, "crowd.yml"), REF)
    subprocess.run([sys.executable, os.path.join(REF, "generate_config.py"), "--config-file", rel],
                   cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    if os.path.isdir(os.path.join(OUT, "config")):
        shutil.rmtree(os.path.join(OUT, "config"))
    shutil.copytree(os.path.join(tmp, "config"), os.path.join(OUT, "config"))
for root, _, files in os.walk(OUT):
    for f in files:
        print(os.path.relpath(os.path.join(root, f), OUT))
