for cfg in "" "EDE_UMMA_GROUP=1" "EDE_UMMA_GROUP=2" "EDE_UMMA_GROUP=3" "EDE_UMMA_GROUP=5"; do
  echo "== $cfg"; env $cfg python tools/run_op.py enb0-hu 64 MFF.up1.conv1+conv2 5 2>&1 | tail -1
  env $cfg python tools/cta_timeline.py enb0-hu 64 MFF.up1.conv1+conv2 2>&1 | grep -E "persistent|issuer waiting for w|producer"
done
