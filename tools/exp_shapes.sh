for shape in 384,2 512,2 256,5 1024,1; do
  echo "== shape $shape"
  VGPU_CACHE_DIR=/tmp/c_$shape VGPU_FUSED_SHAPE=$shape python tools/kbench.py --only orderkey,c4 --iters 5 2>&1 | cut -c1-200
done
