#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/cta_timeline.py enb0-hu 64 D.up4.conv1+conv2 MFF.up1.conv1+conv2 D.up3.conv1+conv2 R.conv0 MFF.conv D.up2.conv1+conv2 D.up4.conv1_2 2>&1 | tail -80
