#!/bin/bash
# Huffman-kernel time against the per-thread chunk size (bits), 224^2 x256 and 1080p x32.
for cb in 256 512 1024 2048 4096 8192; do
  B2P_JPEG_CHUNK_BITS=$cb python benchmarks/jpeg_decode.py --no-pillow --iters 10 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('224 chunk', $cb, d['kernel_ms'], 'host', round(d['host_call_ms'],3), 'exact', d['bit_exact_vs_pillow'])"
done
for cb in 256 1024 4096 8192 16384; do
  B2P_JPEG_CHUNK_BITS=$cb python benchmarks/jpeg_decode.py --no-pillow --iters 4 --size 1080 1920 --batch 32 | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('1080p chunk', $cb, d['kernel_ms'], 'host', round(d['host_call_ms'],3), 'exact', d['bit_exact_vs_pillow'])"
done
