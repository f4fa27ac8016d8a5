"""Probe: does NVML GPM (GPU performance monitoring: DRAM bandwidth utilisation, SM utilisation over an interval)
answer on this box?  If so bench.py can measure DRAM traffic in-run.  Prints what it finds."""
import time
import pynvml as N
import torch

N.nvmlInit()
h = N.nvmlDeviceGetHandleByIndex(0)
try:
    sup = N.nvmlGpmQueryDeviceSupport(h)
    print("gpm supported:", sup.isSupportedDevice)
    s1, s2 = N.nvmlGpmSampleAlloc(), N.nvmlGpmSampleAlloc()
    a = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    b = torch.empty_like(a)
    torch.cuda.synchronize()
    N.nvmlGpmSampleGet(h, s1)
    t0 = time.perf_counter()
    for _ in range(200):
        b.copy_(a)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    N.nvmlGpmSampleGet(h, s2)
    mg = N.c_nvmlGpmMetricsGet_t()
    mg.version = N.NVML_GPM_METRICS_GET_VERSION
    ids = [N.NVML_GPM_METRIC_DRAM_BW_UTIL, N.NVML_GPM_METRIC_SM_UTIL, N.NVML_GPM_METRIC_SM_OCCUPANCY, N.NVML_GPM_METRIC_FP64_UTIL,
           N.NVML_GPM_METRIC_FP32_UTIL, N.NVML_GPM_METRIC_INTEGER_UTIL, N.NVML_GPM_METRIC_GRAPHICS_UTIL]
    mg.numMetrics = len(ids)
    mg.sample1, mg.sample2 = s1, s2
    for i, m in enumerate(ids):
        mg.metrics[i].metricId = m
    N.nvmlGpmMetricsGet(mg)
    for i in range(len(ids)):
        print(ids[i], mg.metrics[i].nvmlReturn, mg.metrics[i].value)
    print("copy GB/s", 200 * 2 * (1 << 30) / dt / 1e9, "interval s", dt)
except Exception as e:
    print("gpm probe failed:", type(e).__name__, e)
