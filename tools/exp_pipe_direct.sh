# one-launch pipeline: tile shape x direct stores (make TUNING=1)
for d in 0 1; do for sh in "64,32,0,256" "64,64,0,256"; do
  echo -n "DIRECT=$d SHAPE=$sh : "; B200_EXP_DIRECT=$d B200_EXP_SHAPE=$sh timeout 25 python tools/exp_pipe.py 2>&1 | tail -1 | sed 's/.*launches\/step, //'
done; done
