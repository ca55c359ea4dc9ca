# what the one-launch pipeline would cost without its TMA load latency (1), without the ready poll (2), without interpolation (4);
# results of flagged runs are wrong by construction -- timing only (make TUNING=1)
for f in ${FLAGS:-0 1 2 3 4 7}; do echo -n "flags=$f : "; B200_EXP_FLAGS=$f timeout 25 python tools/exp_pipe.py 2>&1 | tail -1; done
