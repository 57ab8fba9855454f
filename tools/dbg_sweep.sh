for cfg in "" "EDE_UMMA_SWAP=0" "EDE_UMMA_SWAP=0 EDE_UMMA_CK=32" "EDE_UMMA_SWAP=1 EDE_UMMA_CK=32" "EDE_UMMA_SWAP=1 EDE_UMMA_PACK=2" "EDE_UMMA_SWAP=1 EDE_UMMA_PACK=2 EDE_UMMA_CK=32"; do
  echo "== $cfg"; env $cfg python tools/run_op.py enb0-hu 64 MFF.up1.conv1_2 5 2>&1 | tail -1
  env $cfg python tools/cta_timeline.py enb0-hu 64 MFF.up1.conv1_2 2>&1 | grep -E "persistent"
done
