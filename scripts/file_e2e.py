"""ad-hoc: wall time of the drop-in psearch.exe on files, without / with the index cache"""
import os, subprocess, sys, time, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vntrseek_b200 import synth

nref = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
nread = int(sys.argv[2]) if len(sys.argv) > 2 else 40000
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
exe = os.path.join(root, "vntrseek_b200", "bin", "psearch.exe")
dst = os.path.join(root, "oracle", "_ref", "eucledian.dst")
w = tempfile.mkdtemp()
t = time.time()
refs = synth.make_refs(nref, seed=1)
reads = synth.make_reads(refs, nread, seed=2, rot_frac=0.0)
refs.write_leb36(os.path.join(w, "reference.leb36"))
reads.write_leb36(os.path.join(w, "reads.leb36"))
print("generated %d refs, %d reads in %.1f s (%.0f MB of text)" % (
    nref, nread, time.time() - t, (os.path.getsize(os.path.join(w, "reference.leb36")) + os.path.getsize(os.path.join(w, "reads.leb36"))) / 1e6))
for name, env in (("no cache", {}), ("cache: first run writes it", {"VSK_INDEX_CACHE": "1"}), ("cache: later runs read it", {"VSK_INDEX_CACHE": "1"}),
                  ("cache + 1 load thread", {"VSK_INDEX_CACHE": "1", "VSK_LOAD_THREADS": "1"})):
    t = time.time()
    p = subprocess.run([exe, "reference.leb36", "reads.leb36", dst, "88", "0", "50"], cwd=w, capture_output=True, text=True,
                       env=dict(os.environ, **env))
    dt = time.time() - t
    conf = [l for l in p.stderr.splitlines() if "confirmed" in l]
    print("%-28s rc=%d  %.2f s  %s" % (name, p.returncode, dt, conf[0].strip() if conf else ""))
