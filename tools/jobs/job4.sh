timeout 120 ./tools/probe/tc_probe 2>&1 | tail -8
