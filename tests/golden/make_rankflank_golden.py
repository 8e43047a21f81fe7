"""Makes tests/golden/rankflank_small: synthetic flankalign .map files (make_maps in
tests/test_rankflank.py, seed 5) and the two tables oracle/rankflank_ref.pl -- the reference's
scoring loop in Perl -- prints for them.  Needs perl; the fixture is committed."""
import importlib.util
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
spec = importlib.util.spec_from_file_location("t", os.path.join(ROOT, "tests", "test_rankflank.py"))
t = importlib.util.module_from_spec(spec)
spec.loader.exec_module(t)
out = os.path.join(HERE, "rankflank_small")
shutil.rmtree(out, ignore_errors=True)
t.make_maps(os.path.join(out, "maps"), 5, 3, 40)
subprocess.run(["perl", os.path.join(ROOT, "oracle", "rankflank_ref.pl"), os.path.join(out, "maps"),
                os.path.join(out, "rank.tsv"), os.path.join(out, "map.tsv")], check=True)
