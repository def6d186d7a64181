set -x
mkdir -p gpurun_out
nvidia-smi -L
python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r02a_pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/r02a_bench.json 2> gpurun_out/r02a_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r02a_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['pack_ms'])"
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,sm__inst_executed_pipe_fp64.sum --clock-control none --csv --log-file gpurun_out/r02a_fp64_launches.csv python bench.py --steps 1 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/r02a_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out | tail -5
