python tools/exp_affine.py 2>&1
python tools/exp_pipe.py 2>&1 | tail -1
