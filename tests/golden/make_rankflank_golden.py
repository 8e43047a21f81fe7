"""Makes tests/golden/rankflank_small: synthetic flankalign .map files (make_maps in
tests/test_rankflank.py, seed 5) and the two tables the UNMODIFIED scoring loop of the reference's
perl/run_rankflankmap.pl builds for them (oracle/rankflank_harness.pl evaluates the loop's text, read
from /root/reference where it lies, with the database calls stubbed).  Needs perl and /root/reference;
the fixture is committed.  oracle/rankflank_ref.pl (a restatement, for machines without the
reference tree) must print the same: checked here."""
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
REF_SCRIPT = "/root/reference/perl/run_rankflankmap.pl"
subprocess.run(["perl", os.path.join(ROOT, "oracle", "rankflank_harness.pl"), REF_SCRIPT, os.path.join(out, "maps"),
                os.path.join(out, "rank.tsv"), os.path.join(out, "map.tsv")], check=True)
subprocess.run(["perl", os.path.join(ROOT, "oracle", "rankflank_ref.pl"), os.path.join(out, "maps"),
                "/tmp/_rk_rank.tsv", "/tmp/_rk_map.tsv"], check=True)
assert open("/tmp/_rk_rank.tsv").read() == open(os.path.join(out, "rank.tsv")).read()
assert open("/tmp/_rk_map.tsv").read() == open(os.path.join(out, "map.tsv")).read()
print("rankflank_small:", sum(1 for _ in open(os.path.join(out, "rank.tsv"))), "rank rows")
