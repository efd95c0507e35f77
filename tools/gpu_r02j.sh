#!/bin/bash
# round-2 validation pass: the allocation-free entry point alone in a fresh process, the whole GPU suite, the bench
# lines of every BASELINE.json config, the ncu launch list and one --set full capture of a steady-state step
T=${1:-r02j}
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke(); print(\"smoke ok\")" 2>&1 | tail -1
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/${T}_noalloc_alone.log 2>&1; tail -3 gpurun_out/${T}_noalloc_alone.log
python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/${T}_pytest.log 2>&1; tail -14 gpurun_out/${T}_pytest.log
python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err || tail -5 gpurun_out/${T}_bench.err
python tools/bench_brief.py gpurun_out/${T}_bench.json
python bench.py --mode raster --config cfg0_10k_256_sh0 > gpurun_out/${T}_cfg0.json 2> gpurun_out/${T}_cfg0.err || tail -3 gpurun_out/${T}_cfg0.err
python bench.py --config cfg1_100k_800_sh3 --no-extras > gpurun_out/${T}_cfg1.json 2> gpurun_out/${T}_cfg1.err || tail -3 gpurun_out/${T}_cfg1.err
python bench.py --mode render --config cfg3_4m_1080p_sh3 > gpurun_out/${T}_cfg3.json 2> gpurun_out/${T}_cfg3.err || tail -3 gpurun_out/${T}_cfg3.err
head -c 700 gpurun_out/${T}_cfg0.json; echo; head -c 400 gpurun_out/${T}_cfg1.json; echo; head -c 1200 gpurun_out/${T}_cfg3.json; echo
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
$CMD > gpurun_out/${T}_plain.log 2> gpurun_out/${T}_plain.err || { echo "plain run failed"; tail -20 gpurun_out/${T}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -s 160 -c 60 --csv --log-file gpurun_out/${T}_launches.csv $CMD > gpurun_out/${T}_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'preprocess|depth_key|duplicate|dup_|identify|blend|adam|order_tiles|onesweep|ssim' -s 132 -c 22 -f -o gpurun_out/${T}_full $CMD > gpurun_out/${T}_ncu_full.log 2>&1
tail -2 gpurun_out/${T}_ncu_full.log
