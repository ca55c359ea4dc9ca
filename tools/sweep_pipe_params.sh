# one-launch pipeline: strip rows x frames the strips run ahead x ring size (make TUNING=1)
for sr in ${SRS:-2 4 8}; do for ah in ${AHS:-4 10 14}; do for rm in ${RMS:-24 36}; do
  echo -n "SR=$sr AHEAD=$ah RING_MB=$rm : "; B200_EXP_SR=$sr B200_EXP_AHEAD=$ah B200_EXP_RING_MB=$rm timeout 25 python tools/exp_pipe.py 2>&1 | tail -1 | sed 's/.*launches\/step, //; s/, checksum.*//'
done; done; done
