# C3 standalone (old TMA kernel, RGB smem) with LDS.32 fetch + u32 maps: shapes; px4 paths off for C3
export B200_AFFINE_PX4_C3=0 B200_PIPE_PX4=0
for sh in 0,0,0,0 64,64,0,256 64,32,0,256 32,64,0,256 32,32,0,128; do
  echo "== shape $sh"
  B200_EXP_SHAPE=$sh python tools/exp_affine.py 2>&1 | grep "8UC3"
  B200_EXP_SHAPE=$sh python tools/exp_pipe.py 2>&1 | tail -1
done
