"""Raw output of the FMA-chain microbenchmark behind roofline.peak (nbody_measure_fma_peak),
FP64 and FP32, with the SM clock sampled while it runs.  Writes one JSON object to stdout."""
import json, os, sys, threading, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from parallel_comp_labs_b200 import capi

try:
    import pynvml
    pynvml.nvmlInit()
    h = pynvml.nvmlDeviceGetHandleByIndex(0)
except Exception:
    pynvml = None
out = {"how": "register-resident FMA chains, 8 independent chains x 16 FMA per iteration per thread, "
              "148*8 CTAs x 256 threads; flops = 2 per FMA; CUDA events around one launch",
       "runs": []}
for prec in (64, 32, 64, 32):
    samples, stop = [], threading.Event()

    def poll():
        while not stop.is_set():
            if pynvml:
                samples.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
            stop.wait(0.05)
    th = threading.Thread(target=poll); th.start()
    tf = capi.measure_fma_peak(prec, 1000.0, device=0)
    stop.set(); th.join()
    mhz = sorted(samples)[len(samples) // 2] if samples else None
    lanes = 64 if prec == 64 else 128
    out["runs"].append({"precision": prec, "tflops": tf, "sm_mhz_median": mhz, "ms_target": 1000,
                        "architectural_tflops_at_that_clock":
                            None if not mhz else 148 * lanes * 2 * mhz * 1e6 / 1e12})
if pynvml:
    out["gpu"] = pynvml.nvmlDeviceGetName(h)
    out["sm_max_mhz"] = pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM)
print(json.dumps(out, indent=1))
