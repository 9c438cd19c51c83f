# Live-tracking latency of the shipped library and of every orb-ygz-slam_b200/variants/lib_*.so (kernel experiments, git-ignored).
for lib in "" /root/repo/orb-ygz-slam_b200/variants/lib_*.so; do
  echo -n "variant $(basename ${lib:-shipped}): "
  YGZORB_LIB=$lib python tools/latency_stages.py 640 480 1000 2>&1 | tail -1 | cut -c1-330
done
