"""Run every per-op GPU parity check in its own subprocess (a kernel trap poisons the CUDA context),
with a hard timeout, and write a summary + dumps of failing tensors under gpurun_out/."""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")


def main():
    os.makedirs(OUT, exist_ok=True)
    sys.path.insert(0, ROOT)
    from tests import gpu_checks
    names = sys.argv[1:] or list(gpu_checks.CHECKS)
    env = dict(os.environ, UBR_DUMP_DIR=os.path.join(OUT, "dumps"))
    summary = []
    with open(os.path.join(OUT, "probe.log"), "w") as log:
        for nm in names:
            t0 = time.time()
            try:
                r = subprocess.run(["timeout", "-k", "5", "70", sys.executable, "-m", "tests.gpu_checks", nm], cwd=ROOT,
                                   env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=100)
                rc, out = r.returncode, r.stdout
            except subprocess.TimeoutExpired as e:
                rc, out = -999, (e.stdout or "") + "\nTIMEOUT"
            dt = time.time() - t0
            log.write("===== %s rc=%d (%.1fs)\n%s\n" % (nm, rc, dt, out))
            log.flush()
            tail = "\n".join(out.strip().splitlines()[-12:])
            summary.append((nm, rc))
            print("===== %s rc=%d (%.1fs)\n%s" % (nm, rc, dt, tail), flush=True)
    print("SUMMARY: " + " ".join("%s=%s" % (n, "ok" if rc == 0 else "FAIL(%d)" % rc) for n, rc in summary))


if __name__ == "__main__":
    main()
