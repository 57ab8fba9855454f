for d in 0 1 2 3 5; do echo "== GROUP $d"; EDE_UMMA_GROUP=$d python tools/cta_timeline.py enb0-hu 64 R.conv0 MFF.conv R.conv2 2>&1 | grep -E "persistent|SM clock"; done
