# same-box A/B of two builds of libcpoly.so (tools/ab/libcpoly_old.so, libcpoly_new.so)
for v in old new old new; do
  cp tools/ab/libcpoly_$v.so centerpoly_b200/libcpoly.so
  echo "== $v"
  for cfg in "256 256 64 128 1" "256 64 64 128 1" "512 256 32 64 1"; do echo "  cfg $cfg: $(timeout 120 python tools/fwd_one.py $cfg bf16 1 0.05 2>&1 | tail -1)"; done
  bash tools/frame_ab.sh X=1 | tail -1
done
