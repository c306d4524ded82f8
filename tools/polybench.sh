#!/bin/bash
# The reference's own benchmark/src/poly-benchmark-16k.cpp (BASELINE configs[0]; source untouched, built by
# `make -C oracle polybench`) three ways on this box:
#   ref      the fork as it is, CPU path (single OpenMP thread: the fork's PimData bookkeeping is not thread-safe and the
#            multi-tower cases abort with more)
#   cpu_omp  the hooked build with no handle attached = the reference's CPU path, OpenMP over towers as upstream
#   gpu      the hooked build with FHE_B200_ATTACH=1: every DCRTPoly +, *, SwitchFormat runs on the B200
#   gpu_sync the same with FHE_B200_HOOK_SYNC=1: every operator waits for its kernels (latency per operation; `gpu` is
#            the default enqueue-and-return form, i.e. sustained time per operation with the stream kept busy)
# Writes gpurun_out/polybench_{ref,cpu_omp,gpu,gpu_sync}.json (google-benchmark JSON).
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B=oracle/_ref
ARGS="--benchmark_min_time=${POLYBENCH_MIN_TIME:-0.3} --benchmark_format=json"
OMP_NUM_THREADS=1 $B/poly-benchmark-16k-ref $ARGS > gpurun_out/polybench_ref.json 2> gpurun_out/polybench_ref.err; echo "ref rc=$?"
$B/poly-benchmark-16k-hooked $ARGS > gpurun_out/polybench_cpu_omp.json 2> gpurun_out/polybench_cpu_omp.err; echo "cpu_omp rc=$?"
FHE_B200_ATTACH=1 $B/poly-benchmark-16k-hooked $ARGS > gpurun_out/polybench_gpu.json 2> gpurun_out/polybench_gpu.err; echo "gpu rc=$?"
FHE_B200_ATTACH=1 FHE_B200_HOOK_SYNC=1 $B/poly-benchmark-16k-hooked $ARGS > gpurun_out/polybench_gpu_sync.json 2> gpurun_out/polybench_gpu_sync.err; echo "gpu_sync rc=$?"
python - <<'PY'
import json
runs = {}
for k in ("ref", "cpu_omp", "gpu", "gpu_sync"):
    try:
        runs[k] = {b["name"]: b["real_time"] for b in json.load(open(f"gpurun_out/polybench_{k}.json"))["benchmarks"]}
    except Exception as ex:
        runs[k] = {"error": str(ex)}
names = [n for n in runs.get("ref", {}) if n != "error"]
print(f"{'benchmark':22s} {'ref us':>10s} {'cpu_omp us':>10s} {'gpu us':>10s} {'gpu_sync':>10s} {'ref/gpu':>8s} {'omp/gpu':>8s}")
for n in names:
    r, c, g, gs = runs["ref"].get(n), runs["cpu_omp"].get(n), runs["gpu"].get(n), runs["gpu_sync"].get(n)
    f = lambda v: f"{v:10.1f}" if isinstance(v, (int, float)) else f"{'-':>10s}"
    num = lambda v: isinstance(v, (int, float))
    ratio = f"{r / g:8.1f}" if num(r) and num(g) and g else f"{'':8s}"
    ratio2 = f"{c / g:8.1f}" if num(c) and num(g) and g else ""
    print(f"{n:22s} {f(r)} {f(c)} {f(g)} {f(gs)} {ratio} {ratio2}")
json.dump({"unit": "us per operation (google-benchmark real_time)", "runs": runs}, open("gpurun_out/polybench_summary.json", "w"), indent=1)
PY
