#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s > gpurun_out/r02e_pytest.log 2>&1; tail -6 gpurun_out/r02e_pytest.log
B="python bench.py --no-extras --no-cpu-baseline"
run() { tag=$1; shift; env "$@" $B > gpurun_out/r02e_$tag.json 2> gpurun_out/r02e_$tag.err || tail -3 gpurun_out/r02e_$tag.err; echo "== $tag"; python tools/bench_brief.py gpurun_out/r02e_$tag.json | head -2; }
run base X=1
run sleep50 GS_SORT_SLEEP_NS=50
run sleep200 GS_SORT_SLEEP_NS=200
run dgrid296 GS_DEPTH_SORT_GRID=296 GS_SORT_SLEEP_NS=50
run dgrid148 GS_DEPTH_SORT_GRID=148 GS_SORT_SLEEP_NS=50
run d16 GS_DEPTH_SORT_ITEMS=16 GS_SORT_SLEEP_NS=50
run t16 GS_TILE_SORT_ITEMS=16 GS_SORT_SLEEP_NS=50
run t16d16 GS_TILE_SORT_ITEMS=16 GS_DEPTH_SORT_ITEMS=16 GS_SORT_SLEEP_NS=50
run serial GS_NO_SORT_OVERLAP=1 GS_SORT_SLEEP_NS=50
run grid296 GS_RADIX_GRID=296 GS_SORT_SLEEP_NS=50
