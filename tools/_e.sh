mkdir -p gpurun_out
timeout 600 python tools/bench_embed.py --steps 30 > gpurun_out/r2_embed_bench.json 2> gpurun_out/r2_embed.err; tail -2 gpurun_out/r2_embed.err
timeout 600 python tools/bench_embed.py --steps 20 --fused-cache-rows 1000000 > gpurun_out/r2_embed_fused.json 2>> gpurun_out/r2_embed.err
timeout 900 python tests/measure_decision_flips.py > gpurun_out/r2_decision_flips.json 2>> gpurun_out/r2_embed.err
timeout 600 python tools/bench_api.py > gpurun_out/r2_bench_api.json 2>> gpurun_out/r2_embed.err
timeout 900 python -m pytest tests/test_encoder_gpu.py tests/test_embedder_gpu.py tests/test_dropin_gpu.py -x -q 2>&1 | tail -3
