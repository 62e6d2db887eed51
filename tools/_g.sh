mkdir -p gpurun_out
export BT_AFF=1 BT_XS=1 BT_TC2_LOG=1
for mt in 0 2; do echo "MT=$mt"
for a in "256 256 3 1 1 128 14" "1024 256 1 1 0 128 14" "512 512 3 1 1 128 7" "2048 512 1 1 0 128 7" "128 128 3 1 1 128 28" "512 128 1 1 0 128 28"; do
BT_TC2_MT=$mt timeout 120 python tools/one_layer.py $a 32 3 2>&1 | tail -2 | cut -c1-230
done; done
