# one-launch pipeline: CTAs (of every 8) that prefer strips x strip rows (make TUNING=1)
for sc in ${SC:-0 1 2 3 4}; do for sr in ${SRS:-4 8}; do echo -n "STRIP_CTAS=$sc SR=$sr : "; B200_EXP_STRIP_CTAS=$sc B200_EXP_SR=$sr timeout 25 python tools/exp_pipe.py 2>&1 | tail -1 | cut -c20-; done; done
