#!/bin/bash
# Round-2 evidence capture on one B200 (run under gpurun).  Everything lands in gpurun_out/; the summaries that are
# meant to be judged are copied into profiles/ by hand afterwards.  Numbers printed under ncu are never bench values.
#   $1 = tag (e.g. r02a); $2.. = steps to run: tests bench ref sanitize launches full cli
set -u
TAG=${1:-r02}; shift
STEPS=${*:-tests bench}
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-cpu --cli-targets 0"
for s in $STEPS; do
  case $s in
    tests) timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; tail -5 gpurun_out/${TAG}_pytest_gpu.log ;;
    bench) timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; tail -c 1500 gpurun_out/${TAG}_bench.json; tail -3 gpurun_out/${TAG}_bench.err ;;
    ref) timeout 600 python bench.py --impl reference > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err; tail -c 600 gpurun_out/${TAG}_bench_reference.json ;;
    sanitize)
      for tool in memcheck racecheck initcheck; do
        timeout 900 compute-sanitizer --tool $tool --log-file gpurun_out/${TAG}_sanitizer_$tool.log python profiles/sanitize_batch.py > gpurun_out/${TAG}_sanitizer_$tool.out 2>&1
        echo "$tool rc=$?"; tail -3 gpurun_out/${TAG}_sanitizer_$tool.log; tail -1 gpurun_out/${TAG}_sanitizer_$tool.out
      done ;;
    launches)
      $B > gpurun_out/${TAG}_bench_pre_ncu.log 2>&1 || { echo "bench failed before ncu"; tail -5 gpurun_out/${TAG}_bench_pre_ncu.log; continue; }
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_10000targets.csv $B > gpurun_out/${TAG}_ncu_launch.log 2>&1 ;;
    full)
      for k in ${NCU_KERNELS:-k_merge_best2 k_build_graph2 k_normalize_warp}; do
        timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -f -o gpurun_out/${TAG}_full10k_$k $B > gpurun_out/${TAG}_ncu_full10k_$k.log 2>&1
        echo "ncu $k rc=$?"
        ncu -i gpurun_out/${TAG}_full10k_$k.ncu-rep --page details --csv > gpurun_out/${TAG}_full10k_${k}_details.csv 2>/dev/null
        python profiles/ncu_by_line.py gpurun_out/${TAG}_full10k_$k.ncu-rep $k 45 > gpurun_out/${TAG}_full10k_${k}_by_line.txt 2>/dev/null
      done
      python profiles/make_traffic.py ${TAG} 2 10000 > gpurun_out/${TAG}_traffic.log 2>&1 && cp profiles/traffic.json gpurun_out/${TAG}_traffic.json
      rm -f gpurun_out/${TAG}_full10k_*.ncu-rep ;;
    cli)
      python - <<'PY'
import time, pbdagcon_b200.synth as s
t=time.time(); n=s.write_m5(2,8000,30,0,10000,"/dev/shm/c2_10k.m5"); print("m5 bytes",n,"gen s",round(time.time()-t,1))
PY
      for j in 16 32; do for bt in 256 1024; do t0=$(date +%s.%N); pbdagcon_b200/host/pbdagcon -c 8 -j $j --batch-targets $bt --batch-mb $((bt*3/4)) --stats /dev/shm/c2_10k.m5 > /dev/shm/out_gpu.fa 2> gpurun_out/${TAG}_cli10k_j${j}_b$bt.err; t1=$(date +%s.%N); echo "j=$j batch-targets=$bt wall $(echo "$t1 - $t0" | bc) s"; tail -1 gpurun_out/${TAG}_cli10k_j${j}_b$bt.err; done; done
      rm -f /dev/shm/c2_10k.m5 /dev/shm/out_gpu.fa ;;
  esac
done
ls -la gpurun_out | tail -30
