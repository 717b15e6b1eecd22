#!/bin/bash
# One GPU-box pass that produces the evidence files the roofline and parity claims cite.  Run through gpurun from the repo root:
#   gpurun --timeout 2400 -- 'bash tools/gpu_evidence.sh [tag]'
# Everything lands in gpurun_out/<tag>_*; copy what is to be judged into profiles/.
tag=${1:-r02}
stages=${2:-tests,bench,launches,pipe,full}  # "sanitizer": compute-sanitizer is closed on this pool (profiles/r02a_sanitizer_memcheck_smoke.log)
has() { case ",$stages," in *",$1,"*) return 0;; *) return 1;; esac; }
out=gpurun_out
mkdir -p $out
export PYTHONUNBUFFERED=1
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,memory.total --format=csv > $out/${tag}_gpu.txt 2>&1
nproc >> $out/${tag}_gpu.txt

# 1. parity suite
if has tests; then
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_gputests.log 2>&1
echo "pytest rc=$?" >> $out/${tag}_gputests.log
tail -3 $out/${tag}_gputests.log
fi

# 2. bench (both arms)
if has bench; then
timeout 900 python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err
echo "bench rc=$?"; tail -c 600 $out/${tag}_bench.json
timeout 600 python bench.py --impl reference > $out/${tag}_bench_reference.json 2> $out/${tag}_bench_reference.err
echo "reference rc=$?"
fi

if has launches; then
# 3. ncu launch list of the default workload (no side entries: the list is of the timed step)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $out/${tag}_launches_2p22.csv \
  python bench.py --steps 2 --warmup 1 --no-side --no-cpu-baseline > $out/${tag}_ncu_launches.log 2>&1
echo "ncu launches rc=$?"
fi

if has pipe; then
# 4. pipe counters: the commitment kernel at 2^22 and the fp_mul microbench
PIPE=sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmalite.sum,sm__inst_executed_pipe_alu.sum,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed.sum,sm__cycles_elapsed.avg,sm__cycles_active.avg,smsp__inst_executed_op_integer_pred_on.sum,gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,sm__warps_active.avg.pct_of_peak_sustained_active
ncu --query-metrics 2>/dev/null | grep -i "pipe_fma\|pipe_alu\|pipe_imad\|pipe_int" > $out/${tag}_query_pipe_metrics.txt
timeout 600 ncu --metrics $PIPE --clock-control none -k regex:msm_fixed_kernel -c 2 --csv --log-file $out/${tag}_pipe_msm_fixed_2p22.csv \
  python bench.py --steps 1 --warmup 1 --no-side --no-cpu-baseline > $out/${tag}_ncu_pipe.log 2>&1
echo "ncu pipe rc=$?"
tools/microbench/fpmul_bench > $out/${tag}_fpmul_bench.txt 2>&1
timeout 300 ncu --metrics $PIPE --clock-control none -c 12 --csv --log-file $out/${tag}_pipe_fpmul.csv tools/microbench/fpmul_bench > /dev/null 2>&1
echo "ncu fpmul rc=$?"
fi

if has full; then
# 5. full capture of the dominant kernels at 2^22 (one launch each)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'msm_fixed_kernel|msm_finish_kernel|sc_cubic_eq_kernel|sc_cubic_kernel|sc_quad_half0_kernel|sc_quad_kernel|spmvT3|spmv3' -c 16 \
  -o $out/${tag}_full_2p22 -f python bench.py --steps 1 --warmup 1 --no-side --no-cpu-baseline > $out/${tag}_ncu_full.log 2>&1
echo "ncu full rc=$?"
# the report itself is too large to travel (64 MiB limit of gpurun_out): keep the raw-page CSV and a per-kernel summary
ncu -i $out/${tag}_full_2p22.ncu-rep --page raw --csv > $out/${tag}_full_2p22_raw.csv 2>/dev/null
python tools/ncu_summary.py $out/${tag}_full_2p22.ncu-rep > $out/${tag}_full_2p22_summary.txt 2>&1
ncu -i $out/${tag}_full_2p22.ncu-rep --page source --csv -k regex:msm_fixed_kernel > $out/${tag}_full_2p22_source_msm_fixed.csv 2>/dev/null
sz=$(stat -c %s $out/${tag}_full_2p22.ncu-rep 2>/dev/null || echo 0)
if [ "$sz" -gt 30000000 ]; then rm -f $out/${tag}_full_2p22.ncu-rep; fi
fi

if has sanitizer; then
# 6. compute-sanitizer on the smallest case
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python __graft_entry__.py smoke > $out/${tag}_sanitizer_memcheck_smoke.log 2>&1
echo "memcheck rc=$?"
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python __graft_entry__.py smoke > $out/${tag}_sanitizer_racecheck_smoke.log 2>&1
echo "racecheck rc=$?"
tail -n 5 $out/${tag}_sanitizer_memcheck_smoke.log $out/${tag}_sanitizer_racecheck_smoke.log
fi
du -sh $out
ls -la $out | tail -30
