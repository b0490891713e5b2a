#!/bin/bash
# Development aid: which cluster launch configurations does ncu accept?  (run as ONE ncu invocation over this script)
for cfg in "6 22 256 130000 0" "6 23 256 130000 0" "6 23 256 130000 0 big" "4 24 256 170000 0 big" "4 24 256 170000 1"; do
  build/dsmem_exchange $cfg
done
