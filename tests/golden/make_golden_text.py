#!/usr/bin/env python
"""Text dumps written by the UNMODIFIED reference (BrickMesh::dumpToTextFile / dumpDNPsToTextFile through
oracle/_ref/ref_harness --text), packed as tests/golden/text/<case>.tar.gz: the fixtures of the output path
(doma_b200_dump_text, doma_b200/output.py).  Needs /root/reference and `make -C oracle ref`: this container only."""
import os
import subprocess
import sys
import tarfile
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
HARNESS = os.path.join(HERE, "..", "..", "oracle", "_ref", "ref_harness")
CASES = ["k1_r0_s3", "tr2d_r0_3steps"]   # one steady 3D case, one 2D transient with precursors (IC + 3 steps)

for case in sys.argv[1:] or CASES:
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run([HARNESS, os.path.join(HERE, "cases", case + ".xml"), "1", os.path.join(tmp, "out"), "--no-dump",
                        "--text", os.path.join(tmp, "ref")], check=True, capture_output=True)
        names = sorted(f for f in os.listdir(tmp) if f.startswith("ref") and f.endswith(".txt"))
        with tarfile.open(os.path.join(HERE, "text", case + ".tar.gz"), "w:gz") as tar:
            for n in names:
                tar.add(os.path.join(tmp, n), arcname=n)
        print(case, len(names), "files")
