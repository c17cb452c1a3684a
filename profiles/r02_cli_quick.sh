#!/bin/bash
# Evidence tooling: pbdagcon CLI wall clock on 2000 and 10 000 targets of m5 text (run under gpurun)
python - <<'PY'
import pbdagcon_b200.synth as s
s.write_m5(2,8000,30,0,2000,"/dev/shm/c2_2k.m5"); s.write_m5(2,8000,30,0,10000,"/dev/shm/c2_10k.m5")
PY
for f in c2_2k c2_10k; do for i in 1 2 3; do s=$(date +%s%N); pbdagcon_b200/host/pbdagcon -c 8 -j 16 --stats /dev/shm/$f.m5 2>/tmp/err >/dev/null; e=$(date +%s%N); echo "$f wall $(( (e-s)/1000000 )) ms $(tail -1 /tmp/err | cut -c150-420)"; done; done
rm -f /dev/shm/c2_2k.m5 /dev/shm/c2_10k.m5
