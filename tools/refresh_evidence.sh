# Re-captures the round evidence on a GPU box: GPU tests, ncu launch list + --set full summaries (each after the
# same command ran without ncu), bench lines.  Outputs under gpurun_out/$R/; copy what is to be kept into profiles/.
#   R=r02 sh tools/refresh_evidence.sh
R=${R:-r02}
set -x
mkdir -p gpurun_out/$R
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/$R/pytest_gpu.log 2>&1; tail -2 gpurun_out/$R/pytest_gpu.log
timeout 300 python bench.py --steps 2 --warmup 1 --profile > gpurun_out/$R/profile_plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/$R/launches_bench_profile.csv python bench.py --steps 2 --warmup 1 --profile > gpurun_out/$R/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:argmax_stats_stream -s 3 -c 2 -o /tmp/argmax_u8 -f python bench.py --steps 2 --warmup 1 --profile > gpurun_out/$R/ncu_u8.log 2>&1
python tools/ncu_summary.py /tmp/argmax_u8.ncu-rep gpurun_out/$R/argmax_u8_summary.csv "ncu --set full --clock-control none: python bench.py --steps 2 --warmup 1 --profile (-k regex:argmax_stats_stream -s 3 -c 2), final kernels" > /dev/null
timeout 300 python tools/profile_kernels.py > gpurun_out/$R/profile_kernels_plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -o /tmp/all_kernels -f python tools/profile_kernels.py > gpurun_out/$R/ncu_all.log 2>&1
tail -2 gpurun_out/$R/ncu_all.log
python tools/ncu_summary.py /tmp/all_kernels.ncu-rep gpurun_out/$R/all_kernels_summary.csv "ncu --set full --clock-control none: python tools/profile_kernels.py (every kernel at its benchmark shape, 2 launches each; per-launch times under ncu are cold-cache and serialised)" > /dev/null
timeout 200 python tools/bench_blobs.py > gpurun_out/$R/bench_blobs.log 2>&1
timeout 200 python tools/bench_server.py 2048 64 > gpurun_out/$R/bench_server.log 2>&1
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/$R/bench_reference.json 2> gpurun_out/$R/bench_reference.err
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/$R/bench_final.json 2> gpurun_out/$R/bench_final.err; tail -c 300 gpurun_out/$R/bench_final.err
ls -la gpurun_out/$R
