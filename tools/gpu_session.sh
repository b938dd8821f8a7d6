#!/bin/bash
# One parameterised GPU session (run under gpurun from the repo root).  Every step writes under
# gpurun_out/ with the tag as prefix; copy what should be judged into profiles/.
#
#   tools/gpu_session.sh TAG step [step ...]
#
# steps:
#   tests[:EXPR]        pytest -m gpu -x (optionally -k EXPR); testsall[:EXPR] the same without -x
#   bench[:WL[:FLAGS]]  bench.py --workload WL (default C2) with extra flags ('+' separates flags)
#   ref                 bench.py --impl reference
#   launches[:WL[:FLAGS]]  ncu launch list of the timed steps (gpu__time_duration)
#   ncu:REGEX[:WL[:FLAGS]]  ncu --set full on the kernels matching REGEX, 4 launches
#   san:TOOL:EXPR       compute-sanitizer --tool TOOL on pytest -k EXPR (ONE sanitizer tool per gpurun call)
#   scale:N[:WL[:FLAGS]]  torchrun bench.py at N GPUs, launched as the driver does
#   smoke               __graft_entry__.smoke()
#   wall[:WL]           tools/optimize_wall.py
TAG=${1:?tag}; shift
O=gpurun_out; mkdir -p $O
for step in "$@"; do
    IFS=: read -r kind a b <<< "$step"
    case $kind in
    tests)
        timeout 1500 python -m pytest tests -m gpu -x -q ${a:+-k "$a"} > $O/${TAG}_tests.log 2>&1
        echo "[$TAG] tests exit $?: $(tail -1 $O/${TAG}_tests.log)";;
    testsall)   # no -x: every failure is listed
        timeout 2400 python -m pytest tests -m gpu -q -rf --durations=15 ${a:+-k "$a"} > $O/${TAG}_tests.log 2>&1
        echo "[$TAG] tests exit $?: $(tail -1 $O/${TAG}_tests.log)"
        grep -E "^FAILED|^ERROR" $O/${TAG}_tests.log | cut -c1-260 | head -40;;
    bench)
        wl=${a:-C2}; flags=${b//+/ }
        timeout 1200 python bench.py --workload $wl $flags > $O/${TAG}_bench_${wl}${b:+_${b//[^a-zA-Z0-9]/}}.json \
            2> $O/${TAG}_bench_${wl}.err
        echo "[$TAG] bench $wl $flags exit $?"
        python tools/bench_brief.py $O/${TAG}_bench_${wl}${b:+_${b//[^a-zA-Z0-9]/}}.json;;
    ref)
        timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_ref.err
        echo "[$TAG] reference arm exit $?";;
    launches)
        wl=${a:-C2}; flags=${b//+/ }
        extra="--no-default-sampling --no-lbfgs"; [[ "$flags" == *lbfgs* ]] && extra="--no-default-sampling"
        timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
            --log-file $O/${TAG}_launches_${wl}_raw.csv python bench.py --workload $wl --steps 6 --warmup 3 --no-cpu-baseline \
            $extra $flags > $O/${TAG}_launches_${wl}.log 2>&1
        echo "[$TAG] launch list exit $?"
        python tools/ncu_summary.py launches $O/${TAG}_launches_${wl}_raw.csv > $O/${TAG}_launches_${wl}.csv 2>&1
        tail -12 $O/${TAG}_launches_${wl}.csv;;
    ncu)
        IFS=: read -r kind a wl flags <<< "$step"
        wl=${wl:-C2}; flags=${flags//+/ }
        timeout 900 ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"$a" -c 4 \
            -o $O/${TAG}_${wl}_full python bench.py --workload $wl --steps 1 --warmup 3 --no-cpu-baseline \
            --no-default-sampling --no-lbfgs $flags > $O/${TAG}_ncu_${wl}.log 2>&1
        echo "[$TAG] ncu --set full exit $?"
        python tools/ncu_summary.py details $O/${TAG}_${wl}_full.ncu-rep > $O/${TAG}_${wl}_ncu_details.txt 2>&1
        grep -E "^==|gpu__time_duration|dram__bytes|issue_active" $O/${TAG}_${wl}_ncu_details.txt | head -40;;
    san)
        timeout 1500 compute-sanitizer --tool $a --error-exitcode 9 python -m pytest tests -m gpu -x -q -k "$b" \
            > $O/${TAG}_sanitizer_${a}.log 2>&1
        echo "[$TAG] compute-sanitizer $a exit $?: $(grep -E 'ERROR SUMMARY|passed|failed' $O/${TAG}_sanitizer_${a}.log | tail -2 | tr '\n' ' ')";;
    scale)
        IFS=: read -r kind a wl flags <<< "$step"
        wl=${wl:-C2}; flags=${flags//+/ }
        timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $a --master-addr 127.0.0.1 --master-port 29511 \
            bench.py --gpus $a --workload $wl --steps 10 --warmup 3 $flags > $O/${TAG}_bench_${wl}_${a}gpu.json 2> $O/${TAG}_bench_${wl}_${a}gpu.err
        echo "[$TAG] scale $a exit $?"
        python tools/bench_brief.py $O/${TAG}_bench_${wl}_${a}gpu.json;;
    smoke)
        timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1
        echo "[$TAG] smoke exit $?: $(tail -1 $O/${TAG}_smoke.log)";;
    wall)
        timeout 900 python tools/optimize_wall.py ${a:-C2} > $O/${TAG}_optimize_wall_${a:-C2}.jsonl 2> $O/${TAG}_wall.err
        echo "[$TAG] optimize wall exit $?: $(tail -1 $O/${TAG}_optimize_wall_${a:-C2}.jsonl | cut -c1-300)";;
    *) echo "unknown step $step";;
    esac
done
