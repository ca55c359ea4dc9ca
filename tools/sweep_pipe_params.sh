for sr in 2 4 8; do for ah in 4 8 10 14; do for rm in 16 24 36; do
  echo -n "SR=$sr AHEAD=$ah RING_MB=$rm : "; B200_EXP_SR=$sr B200_EXP_AHEAD=$ah B200_EXP_RING_MB=$rm timeout 60 python tools/exp_pipe.py 2>&1 | tail -1 | sed 's/.*launches\/step, //; s/, checksum.*//'
done; done; done
