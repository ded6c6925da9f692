#!/bin/bash
# build libpcp_cuda.so from anywhere; prints errors or the register line of the kernels named in $@
cd /root/repo && python -m pointcloudprocess_b200.build -v > /tmp/build.log 2>&1 || { grep -E "error|Error" -A3 /tmp/build.log | head -40; exit 1; }
for k in "$@"; do grep -A2 "Compiling entry function.*$k" /tmp/build.log | grep -E "Used|spill" ; done
echo built
