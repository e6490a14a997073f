for lib in libvaekit.so libvaekit_tcB.so libvaekit_tcC.so; do
  echo "== $lib"
  VAEKIT_LIB=$PWD/pytorch-vae_b200/vaekit/_lib/$lib timeout 300 python scripts/time_ops.py tc 2>&1 | tail -2
  VAEKIT_LIB=$PWD/pytorch-vae_b200/vaekit/_lib/$lib timeout 300 python -m pytest tests/test_gpu_betatc.py -x -q 2>&1 | tail -1
done
