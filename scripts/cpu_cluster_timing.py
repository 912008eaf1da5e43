"""Timed CPU run of BASELINE config 2: the reference's default cluster (1 ps + 2 workers, sync, batch 128 per worker,
CNN, Adam lr 0.01, keep_prob 0.75) as three OS processes over gloo on this machine, started through the role
scripts like /root/reference/mnist_python_{m,w1,w2}.py.  Prints the workers' own 'Training elapsed time' lines and
the aggregate images/s."""
import os
import re
import socket
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
common = ["--ps_hosts=127.0.0.1:%d" % port, "--worker_hosts=127.0.0.1:%d,127.0.0.1:%d" % (port + 1, port + 2),
          "--backend=cpu", "--train_steps=%d" % steps, "--validation_batches=1", "--validation_batch_size=200",
          "--step_timeout_secs=300"]
env = dict(os.environ, PYTHONPATH=ROOT, OMP_NUM_THREADS=str(max(1, (os.cpu_count() or 2) // 2)))
procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, sc)] + common, stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, text=True, env=env, cwd=ROOT)
         for sc in ("mnist_python_m.py", "mnist_python_w1.py", "mnist_python_w2.py")]
outs = [p.communicate(timeout=3600)[0] for p in procs]
for p, o in zip(procs, outs):
    if p.returncode != 0:
        print(o[-2000:])
        sys.exit(1)
secs = [float(re.search(r"Training elapsed time: ([0-9.]+) s", o).group(1)) for o in outs[1:]]
print("workers' elapsed: %s s for %d global steps" % (secs, steps))
print("aggregate: %.1f images/s (2 workers x 128 per global step)" % (2 * 128 * steps / max(secs)))
