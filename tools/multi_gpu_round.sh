N=$1
for comm in p2p nccl; do CGB200_COMM=$comm CGB200_REQUIRE_P2P=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N tests/dist_check.py > gpurun_out/r02_dist_check_n${N}_$comm.log 2>&1; echo "dist_check N=$N $comm rc=$? pass=$(grep -c PASS gpurun_out/r02_dist_check_n${N}_$comm.log) fail=$(grep -c FAIL gpurun_out/r02_dist_check_n${N}_$comm.log)"; done
timeout 700 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err; echo "bench N=$N rc=$?"; python - <<PY
import json
d=json.loads(open("gpurun_out/r02_bench_n$N.json").read().strip().splitlines()[-1])
print(json.dumps(d["summary"]))
print(json.dumps(d.get("comm",{}))[:1500])
PY
