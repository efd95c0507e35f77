#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s > gpurun_out/r02i_pytest.log 2>&1; tail -6 gpurun_out/r02i_pytest.log
python bench.py > gpurun_out/r02i_bench.json 2> gpurun_out/r02i_bench.err || tail -5 gpurun_out/r02i_bench.err
python tools/bench_brief.py gpurun_out/r02i_bench.json
python bench.py --mode raster --config cfg0_10k_256_sh0 > gpurun_out/r02i_cfg0.json 2> gpurun_out/r02i_cfg0.err || tail -3 gpurun_out/r02i_cfg0.err
python bench.py --config cfg1_100k_800_sh3 --no-extras > gpurun_out/r02i_cfg1.json 2> gpurun_out/r02i_cfg1.err || tail -3 gpurun_out/r02i_cfg1.err
python bench.py --mode render --config cfg3_4m_1080p_sh3 > gpurun_out/r02i_cfg3.json 2> gpurun_out/r02i_cfg3.err || tail -3 gpurun_out/r02i_cfg3.err
python bench.py --batch-views 64 --steps 3 --warmup 3 > gpurun_out/r02i_cfg4.json 2> gpurun_out/r02i_cfg4.err || tail -3 gpurun_out/r02i_cfg4.err
head -c 700 gpurun_out/r02i_cfg0.json; echo; head -c 400 gpurun_out/r02i_cfg1.json; echo; head -c 900 gpurun_out/r02i_cfg3.json; echo; head -c 900 gpurun_out/r02i_cfg4.json; echo
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
$CMD > gpurun_out/r02i_plain.log 2> gpurun_out/r02i_plain.err || { echo "plain run failed"; tail -20 gpurun_out/r02i_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 160 -c 60 --csv --log-file gpurun_out/r02i_launches.csv $CMD > gpurun_out/r02i_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'preprocess|depth_key|duplicate|dup_|identify|blend|adam|order_tiles|onesweep|ssim' -s 132 -c 22 -f -o gpurun_out/r02i_full $CMD > gpurun_out/r02i_ncu_full.log 2>&1
tail -2 gpurun_out/r02i_ncu_full.log
