mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) > gpurun_out/t.log; cat gpurun_out/t.log
export BT_AFF=1 BT_XS=1
for rw in 0 3 2; do
  echo "RW=$rw"
  BT_TC2_RW=$rw timeout 120 python tools/one_layer.py 256 256 3 1 1 128 14 32 3 2>&1 | tail -1
  BT_TC2_RW=$rw timeout 120 python tools/one_layer.py 1024 256 1 1 0 128 14 32 3 2>&1 | tail -1
  BT_TC2_RW=$rw timeout 120 python tools/one_layer.py 512 512 3 1 1 128 7 32 3 2>&1 | tail -1
  BT_TC2_RW=$rw timeout 120 python tools/one_layer.py 2048 512 1 1 0 128 7 32 3 2>&1 | tail -1
  BT_TC2_RW=$rw BT_RES=1 timeout 120 python tools/one_layer.py 512 2048 1 1 0 128 7 32 3 2>&1 | tail -1
  BT_TC2_RW=$rw timeout 120 python tools/one_layer.py 128 128 3 1 1 128 28 32 3 2>&1 | tail -1
done
unset BT_AFF BT_XS
timeout 300 python tools/layer_table.py 32 > gpurun_out/layer_table_mc32_b.txt 2>&1; tail -32 gpurun_out/layer_table_mc32_b.txt
