# warp-specialised pipeline (make TUNING=1): staged bytes per source row and stage x strip rows x direct stores
for d in ${DS:-0 1}; do for sr in ${SRS:-4}; do for st in ${STS:-1920 2880 3840 5760}; do
  echo -n "PIPE2 DIRECT=$d STAGE_ROW=$st SR=$sr : "; B200_EXP_PIPE2=1 B200_EXP_DIRECT=$d B200_EXP_STAGE_ROW=$st B200_EXP_SR=$sr timeout 25 python tools/exp_pipe.py 2>&1 | tail -1 | sed 's/.*launches\/step, //'
done; done; done
