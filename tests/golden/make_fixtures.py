"""Regenerates the fixtures in this directory from the read-only reference checkout.

The files are DATA (the reference's example filterbank and the golden output its authors
committed), not source code:
  tutorial.fil          example_data/tutorial.fil            (2-bit, 64 chan, 187520 samples)
  candidates.peasoup    example_output/candidates.peasoup    (10 candidates, 1032 detections)
  overview.xml          example_output/overview.xml          (parameters, DM list, accel list)
  default_zaplist.txt, 47tuc.zaplist   misc/                 (birdie lists: "freq width" lines)
Run here (the GPU box has no /root/reference):  python tests/golden/make_fixtures.py
"""
import os
import shutil

REF = os.environ.get("PEASOUP_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
for rel in ["example_data/tutorial.fil", "example_output/candidates.peasoup", "example_output/overview.xml",
            "misc/default_zaplist.txt", "misc/47tuc.zaplist"]:
    shutil.copyfile(os.path.join(REF, rel), os.path.join(HERE, os.path.basename(rel)))
    print("copied", rel)
