for cfg in "EDE_UMMA_TAIL=1" "EDE_UMMA_TAIL=0" "EDE_UMMA_TAIL=1 EDE_UMMA_GROUP=1" "EDE_UMMA_TAIL=1 EDE_UMMA_GROUP=3"; do
  echo "== $cfg"; env $cfg python tools/run_op.py enb0-hu 64 R.conv0 5 2>&1 | tail -1
  env $cfg python tools/cta_timeline.py enb0-hu 64 R.conv0 2>&1 | grep -E "persistent|issuer waiting|producer|SM clock|epilogue waiting"
done
