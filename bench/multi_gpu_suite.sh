#!/usr/bin/env bash
# Everything measured on N GPUs of one box, each step under its own timeout:   bash bench/multi_gpu_suite.sh N [what...]
#   what: test sweep sweep16 bw bench ref   (default: all)
# Outputs land in gpurun_out/ (bench_n<N>_{ours,reference}.json, allreduce_sweep_w<N>*.json, p2p_bandwidth_w<N>.json).
set -u
N="${1:-2}"; shift || true
WHAT="${*:-test bw sweep sweep16 bench ref}"
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
port() { python - <<'PY'
import socket
s = socket.socket(); s.bind(("127.0.0.1", 0)); print(s.getsockname()[1]); s.close()
PY
}
for w in $WHAT; do
  case $w in
    test)    timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_multi_n$N.log 2>&1; echo "test rc=$?"; tail -3 gpurun_out/pytest_multi_n$N.log ;;
    bw)      timeout 300 $TR --master-port $(port) bench/p2p_bandwidth.py > gpurun_out/p2p_bw_n$N.log 2>&1; echo "bw rc=$?"; grep -E "GB/s" gpurun_out/p2p_bw_n$N.log | tail -20 ;;
    sweep)   SWEEP_BLOCKS="${SWEEP_BLOCKS:-16,32,64}" timeout 600 $TR --master-port $(port) bench/allreduce_sweep.py > gpurun_out/sweep_n$N.log 2>&1; echo "sweep rc=$?"; grep " B " gpurun_out/sweep_n$N.log | tail -30 ;;
    sweep16) SWEEP_DTYPE=bf16 SWEEP_BLOCKS="${SWEEP_BLOCKS:-32,64}" timeout 600 $TR --master-port $(port) bench/allreduce_sweep.py > gpurun_out/sweep16_n$N.log 2>&1; echo "sweep16 rc=$?"; grep " B " gpurun_out/sweep16_n$N.log | tail -8 ;;
    bench)   timeout 900 $TR --master-port $(port) bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n${N}_ours.json 2> gpurun_out/bench_n${N}_ours.err; echo "bench rc=$?"; tail -c 300 gpurun_out/bench_n${N}_ours.err ;;
    ref)     timeout 900 $TR --master-port $(port) bench.py --impl reference --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n${N}_reference.json 2> gpurun_out/bench_n${N}_reference.err; echo "ref rc=$?" ;;
  esac
done
python - "$N" <<'PY'
import json, sys
n = sys.argv[1]
for arm in ("ours", "reference"):
    try:
        d = json.loads(open(f"gpurun_out/bench_n{n}_{arm}.json").read().strip().splitlines()[-1])
    except Exception as e:
        print(arm, "no line:", e); continue
    print(arm, "resnet50", round(d["value"]), "img/s", round(d["ms_per_step"], 2), "ms e2e", round(d["e2e"]["value"]))
    for s in d.get("secondary", []):
        print("   ", s["config"]["model"], s["config"]["image_size"], round(s.get("value", 0), 1), "img/s", round(s.get("ms_per_step", 0), 3), "ms", s.get("error", ""))
PY
