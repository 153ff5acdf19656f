"""Run every test function of a pytest file in its own process (a CUDA trap in one kernel variant poisons the
context of that process only).  Usage: python tools/run_isolated.py tests/test_gpu_ops.py [-m gpu] -> gpurun_out/iso_*.log"""
import os
import re
import subprocess
import sys

path = sys.argv[1]
extra = sys.argv[2:]
os.makedirs("gpurun_out", exist_ok=True)
names = re.findall(r"^def (test_\w+)", open(path).read(), flags=re.M)
summary = []
for n in names:
    log = os.path.join("gpurun_out", f"iso_{os.path.basename(path)[:-3]}_{n}.log")
    try:
        r = subprocess.run([sys.executable, "-m", "pytest", path, "-k", n, "-q", "-x", "--tb=short", "-p", "no:cacheprovider"] + extra,
                           capture_output=True, text=True, timeout=600)
        out, rc = r.stdout + r.stderr, r.returncode
    except subprocess.TimeoutExpired as e:
        out, rc = (e.stdout or b"").decode(errors="replace") + "\nTIMEOUT", -9
    open(log, "w").write(out)
    tail = [l for l in out.strip().splitlines() if l.strip()][-1:] or [""]
    summary.append(f"{n}: rc={rc} :: {tail[0]}")
    print(summary[-1], flush=True)
open(os.path.join("gpurun_out", f"iso_{os.path.basename(path)[:-3]}_summary.txt"), "w").write("\n".join(summary) + "\n")
