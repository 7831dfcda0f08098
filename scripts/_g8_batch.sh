set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/bench_g8_v12.json 2> gpurun_out/bench_g8_v12.err; cat gpurun_out/bench_g8_v12.json | cut -c1-300
timeout 900 $TR scripts/train_epochs.py --latent-bits 24 --epochs 2 --precision tc32 --algorithm pairs > gpurun_out/train_g8_pairs_v12.log 2> gpurun_out/train_g8_pairs_v12.err; tail -1 gpurun_out/train_g8_pairs_v12.log
timeout 600 $TR scripts/train_epochs.py --latent-bits 24 --epochs 40 --precision fp64 --algorithm modeproduct > gpurun_out/train_g8_mp_v12.log 2> gpurun_out/train_g8_mp_v12.err; tail -1 gpurun_out/train_g8_mp_v12.log
