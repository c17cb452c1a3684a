set -x
python -m pytest tests/test_gpu_golden.py -x -q > gpurun_out/pytest_gpu_golden.log 2>&1; tail -15 gpurun_out/pytest_gpu_golden.log
python - <<'PY'
import time, pbdagcon_b200.synth as s
t=time.time(); n=s.write_m5(2,8000,30,0,2000,"/dev/shm/c2.m5"); print("m5 bytes",n,"gen s",time.time()-t)
PY
for j in 4 16; do sync; time pbdagcon_b200/host/pbdagcon -c 8 -j $j --stats /dev/shm/c2.m5 > /dev/shm/out_gpu.fa 2> gpurun_out/cli_gpu_j$j.err; tail -25 gpurun_out/cli_gpu_j$j.err | grep -E "mbp_per_s"; done
time oracle/_ref/pbdagcon_ref -j 16 -c 8 /dev/shm/c2.m5 > /dev/shm/out_ref.fa 2> gpurun_out/cli_ref.err; true
python - <<'PY'
def canon(p):
    L=open(p).read().split("\n"); return sorted(L[i]+L[i+1] for i in range(0,len(L)-1,2))
a,b=canon("/dev/shm/out_gpu.fa"),canon("/dev/shm/out_ref.fa")
print("records",len(a),len(b),"identical",a==b, "bases", sum(len(x) for x in open("/dev/shm/out_gpu.fa").read().split("\n")[1::2]))
PY
