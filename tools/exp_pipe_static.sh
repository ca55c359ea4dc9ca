# upper bound of role specialisation: fixed roles + static work lists, no queue (make TUNING=1; idle GPU only)
for ns in ${NS:-1 2 3}; do for sr in ${SRS:-4 8}; do echo -n "STATIC=$ns SR=$sr : "; B200_EXP_STATIC=$ns B200_EXP_SR=$sr timeout 25 python tools/exp_pipe.py 2>&1 | tail -1 | cut -c20-; done; done
