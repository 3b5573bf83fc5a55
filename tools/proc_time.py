"""Development helper: wall time of palace_b200/bin/eref as a process (EREF_TIMING=1) on a BASELINE configs[0]-sized
sample, three runs, on a GPU box.  usage: python tools/proc_time.py"""
import os, subprocess, sys, tempfile, time
sys.path.insert(0, ".")
from palace_b200 import synth
d = tempfile.mkdtemp()
db = synth.make_db(1000, 50_000_000, 1000)
f1, f2, _ = synth.make_reads(db, 5_000_000, 2000)
db.text.tofile(d + "/db.fa"); f1.tofile(d + "/r1.fq"); f2.tofile(d + "/r2.fq")
env = dict(os.environ, EREF_GPUS="1", EREF_TIMING="1", EREF_CODER_SEED="1")
for i in range(3):
    t0 = time.perf_counter()
    r = subprocess.run(["palace_b200/bin/eref", d + "/r1.fq", d + "/r2.fq", d + "/db.fa", d + "/iv.txt", "0.9", "0.85", "16"], capture_output=True, env=env)
    print(f"run {i}: {time.perf_counter() - t0:.3f} s rc={r.returncode} lines={r.stdout.count(b'\n')}")
    print(r.stderr.decode().splitlines()[0] if r.stderr else "")
t0 = time.perf_counter(); subprocess.run(["python", "-c", "import ctypes; ctypes.CDLL('libcudart.so.12'); ctypes.CDLL('libnccl.so.2')"]); print("python + dlopen cudart+nccl", round(time.perf_counter() - t0, 3))
