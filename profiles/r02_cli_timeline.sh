#!/bin/bash
# Evidence tooling: verbose timeline of the pbdagcon CLI on 10 000 targets of m5 text (run under gpurun)
python - <<'PY'
import time, pbdagcon_b200.synth as s
t=time.time(); n=s.write_m5(2,8000,30,0,10000,"/dev/shm/c2_10k.m5"); print("m5 bytes",n,"gen s",round(time.time()-t,1))
PY
for bt in 256 512 1024; do
  echo "== batch-targets $bt"
  pbdagcon_b200/host/pbdagcon -v -c 8 -j 16 --batch-targets $bt --batch-mb $((bt*3/4)) /dev/shm/c2_10k.m5 2>&1 >/dev/null | grep -E "ready|submit batch (0|1|2|3|8|20) |collected batch (0|3|8|20|39) |targets\"" | cut -c1-200
  pbdagcon_b200/host/pbdagcon -c 8 -j 16 --stats --batch-targets $bt --batch-mb $((bt*3/4)) /dev/shm/c2_10k.m5 2>&1 >/dev/null | tail -1 | cut -c100-400
done
rm -f /dev/shm/c2_10k.m5
