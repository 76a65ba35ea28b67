nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active,temperature.gpu --format=csv,noheader -lms 50 > gpurun_out/knn_clocks.csv &
SMI=$!
sleep 0.5
timeout 100 python tools/knn_sharded.py --nq 1000000 --nt 1000000 --engine fp4 --steps 20 --warmup 2 2>&1 | tail -1 | cut -c95-130
kill $SMI
sleep 0.2
nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active,temperature.gpu --format=csv,noheader -lms 50 > gpurun_out/mma_clocks.csv &
SMI=$!
sleep 0.3
tools/mma_rate_probe | grep tcmp
kill $SMI
