set -e
cd /root/repo
python - <<'PY'
import sys, os, time, subprocess, tempfile
sys.path.insert(0, '.')
from bramble_b200 import workloads
import bench
W = workloads.build('c2', 1000, 300000)
rb = bench.RefBench(W, {"small": 2000, "cli": 300000})
import __graft_entry__ as g
cli = g.build_cli()
for env in ({}, {"BRB_INFLATE_THREADS": "12"}, {"BRB_DEVICE_INFLATE": "1"}):
    for rep in range(2):
        wd = tempfile.mkdtemp(dir=rb.d)
        e = dict(os.environ, BRB_CLI_TIMING="1", **env)
        t0 = time.time()
        r = subprocess.run([cli, rb.d + "/cli.bam", "-G", rb.d + "/a.gtf", "-o", wd + "/out.bam", "--quiet", "--fr", "-p", "16", "--level", "1"], cwd=wd, env=e, capture_output=True, text=True)
        print(env, rep, round(time.time() - t0, 2), r.returncode, r.stderr.strip()[-2400:])
PY
