#!/bin/bash
# Round-2 profile set (run on a B200 through gpurun; everything lands in gpurun_out/, summaries are made in the build
# container with tools/ncu_summary.py / tools/ncu_launches.py / tools/traffic_from_ncu.py and committed under profiles/).
cd "$(dirname "$0")/.."
O=gpurun_out
# 1. the bench line itself (never under a profiler)
python bench.py > $O/r02_bench_n1.json 2> $O/r02_bench_n1.err; echo "bench rc=$?"
# 2. launch list of the same command (cold-cache, serialised: shares, not absolutes)
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/r02_launches.csv \
    python bench.py --steps 2 --warmup 3 --headline-only > $O/ncu_launches.log 2>&1; echo "launch list rc=$?"
# 3. full captures of one eager step (after a warm-up step): every GEMM launch, then every other kernel of ours
python tools/step_once.py > $O/step_once.log 2>&1; tail -1 $O/step_once.log
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -s 48 -c 48 -o $O/r02_gemm_full \
    python tools/step_once.py > $O/ncu_gemm.log 2>&1; echo "gemm capture rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"attn_|lstm_|conv0_|rownorm_|nce_|adam_kernel" -c 60 \
    -o $O/r02_other_full python tools/step_once.py --steps 2 > $O/ncu_other.log 2>&1; echo "other capture rc=$?"
# 4. the long-window attention shape (S = 244, dk = 64) and the default one, single launches
ncu --set full --clock-control none --import-source on -k regex:"attn_" -o $O/r02_attn_shapes \
    python tools/attn_bench.py --once > $O/ncu_attn.log 2>&1; echo "attn capture rc=$?"
# the reports stay on the box (gpurun merges at most 64 MiB back): summaries and raw metric tables are made here
for r in r02_gemm_full r02_other_full r02_attn_shapes; do
  python tools/ncu_summary.py $O/$r.ncu-rep > $O/${r}_summary.txt 2>&1
  ncu -i $O/$r.ncu-rep --page raw --csv 2>/dev/null | gzip > $O/${r}_raw.csv.gz
done
python tools/traffic_from_ncu.py $O/r02_gemm_full.ncu-rep $O/r02_other_full.ncu-rep > $O/r02_traffic.json 2> $O/traffic.err
python tools/ncu_launches.py $O/r02_launches.csv "ncu --metrics gpu__time_duration.sum --clock-control none -c 900 python bench.py --steps 2 --warmup 3 --headline-only" > $O/r02_ncu_launches.txt 2>&1
# source-level stall samples of three representative GEMM launches (conv1 fused-norm, the heads' FFN1, a split-K wgrad)
for k in 0 11 15; do
  ncu -i $O/r02_gemm_full.ncu-rep --page source --csv --print-kernel-base function --launch-skip $k --launch-count 1 2>/dev/null | gzip > $O/r02_gemm_source_launch$k.csv.gz
done
mkdir -p /tmp/ncu_reports && mv $O/*.ncu-rep /tmp/ncu_reports/ && rm -f $O/r02_launches.csv.bak
# 5. timelines of one graph replay (CUPTI), launch marks inside a real replay, per-launch event profile
python tools/timeline.py > $O/r02_timeline.txt 2>&1
python tools/timeline_dp.py > $O/r02_timeline_dp_n1.txt 2>/dev/null
python tools/graph_marks.py > $O/r02_graph_marks.txt 2>&1
python tools/step_profile.py > $O/r02_step_profile.txt 2>&1
python tools/step_profile.py hiddenEncoder=512 hiddenGar=512 nLevelsGRU=2 sizeWindow=40960 > $O/r02_step_profile_cfg4.txt 2>&1
python tools/step_profile.py arMode=transformer > $O/r02_step_profile_cfg3.txt 2>&1
python tools/gemm_phase_probe.py > $O/r02_gemm_phases.txt 2>&1
python tools/attn_bench.py > $O/r02_attn_bench.txt 2>&1
python tools/lstm_bench.py > $O/r02_lstm_bench.txt 2>&1
du -sh $O
