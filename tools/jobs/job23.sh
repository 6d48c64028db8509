timeout 120 ./build/tg_probe 2368 3
timeout 120 ./build/tg_probe 2368 1
