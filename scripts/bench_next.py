"""Times the next-row kernels on the GPU and the reference formulas on the host cores.
python scripts/bench_next.py > profiles/r01_next_rows.json"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from pointdsc_b200 import _cabi, matching, metrics
from oracle.gen_golden_next import descriptors, metric_inputs
from oracle import restate_next


def gpu_ms(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


out = {}
# N2: LiDAR front end, 12 000 keypoints, 32-dim descriptors (datasets/LidarFeatureExtractor.py), 16 pairs per call
P, N, D = 16, 12000, 32
F0, F1 = descriptors(1, N, N, D)
A = F0[None].repeat(P, 1, 1).cuda()
B = F1[None].repeat(P, 1, 1).cuda()


def gpu_match():
    iab = _cabi.nn_match(A, B, metric=_cabi.MATCH_DOT)[0]
    iba = _cabi.nn_match(B, A, metric=_cabi.MATCH_DOT)[0]
    return _cabi.mutual_corr(iab, iba, N)


ms = gpu_ms(gpu_match, reps=5)
t0 = time.perf_counter()
restate_next.lidar_mutual_corr(F0.numpy(), F1.numpy())
cpu_s = time.perf_counter() - t0
flops = 2.0 * 2 * N * N * D * P                       # both directions
out["mutual_nn_correspondences"] = {
    "workload": f"{P} pairs x {N} x {N} x {D}-dim descriptors, mutual", "gpu_ms_per_call": ms, "pairs_per_s": P / ms * 1e3,
    "fp32_tflops": flops / ms / 1e9, "cpu_reference_s_per_pair": cpu_s, "cpu_threads": torch.get_num_threads(),
    "speedup_vs_cpu": cpu_s / (ms / 1e3 / P)}
# N3: metrics for 1024 pairs of 5000 correspondences
Pm, Nm = 1024, 5000
src, tgt, est, gt, logits, labels = metric_inputs(5, 4, Nm)
rep = lambda x: x.repeat(Pm // 4, *([1] * (x.dim() - 1))).cuda().contiguous()
src, tgt, est, gt, logits, labels = map(rep, (src, tgt, est, gt, logits, labels))
ms_r = gpu_ms(lambda: metrics.registration_errors(est, gt, src, tgt))
ms_c = gpu_ms(lambda: _cabi.classification_counts(logits, labels))
t0 = time.perf_counter()
restate_next.transformation_metrics(est[:16].cpu(), gt[:16].cpu(), src[:16].cpu(), tgt[:16].cpu())
cpu_r = (time.perf_counter() - t0) / 16
out["registration_metrics"] = {"workload": f"{Pm} pairs x {Nm} correspondences", "gpu_ms_per_call": ms_r,
                               "effective_GBps": Pm * Nm * 24 / ms_r / 1e6, "cpu_port_s_per_pair": cpu_r}
out["classification_counts"] = {"workload": f"{Pm} pairs x {Nm} correspondences", "gpu_ms_per_call": ms_c,
                                "effective_GBps": Pm * Nm * 8 / ms_c / 1e6}
print(json.dumps(out, indent=1))
