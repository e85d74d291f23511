"""Build and run tools/micro/fp64_peak.cu on the box; write the DFMA/DMMA peaks + clocks as JSON (profiles/r02/fp64_peak.json)."""
import json, os, re, subprocess, sys, time
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = os.path.join(root, "tools", "micro", "fp64_peak.cu")
exe = os.path.join(root, "tools", "micro", "fp64_peak")
subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-o", exe, src])
q = ["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,clocks.mem,power.draw,clocks_throttle_reasons.active", "--format=csv,noheader"]
p = subprocess.Popen([exe], stdout=subprocess.PIPE, text=True)
clk = []
while p.poll() is None:
    try:
        clk.append(subprocess.check_output(q, text=True).strip().splitlines()[0])
    except Exception as e:  # noqa
        pass
    time.sleep(0.05)
out = p.stdout.read()
dfma = [float(x) for x in re.findall(r"DFMA.*?: ([0-9.]+) TFLOP/s", out)]
dmma = [float(x) for x in re.findall(r"DMMA.*?: ([0-9.]+) TFLOP/s", out)]
sm = sorted(int(c.split(",")[0].split()[0]) for c in clk) if clk else [0]
rec = {"dfma_tflops": max(dfma), "dmma_tflops": max(dmma), "dfma_all": dfma, "dmma_all": dmma,
       "sm_mhz_median": sm[len(sm) // 2], "sm_mhz_max_seen": sm[-1], "clock_samples": clk[:: max(1, len(clk) // 8)],
       "gpu": out.strip().splitlines()[-1], "how": "tools/micro/fp64_peak.cu: 16 independent DFMA chains / 8 independent m8n8k4 DMMA chains per thread, CUDA events, best launch shape",
       "when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()), "raw": out}
dst = sys.argv[1] if len(sys.argv) > 1 else os.path.join(root, "gpurun_out", "fp64_peak.json")
json.dump(rec, open(dst, "w"), indent=1)
print(json.dumps({k: rec[k] for k in ("dfma_tflops", "dmma_tflops", "sm_mhz_median")}))
