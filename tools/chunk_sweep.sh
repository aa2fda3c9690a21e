for C in 0 512 256 128 64; do echo "== ORBX_DEVICE_CHUNK=$C"; ORBX_DEVICE_CHUNK=$C timeout 120 python bench.py --steps 20 --warmup 3 --value-only 2>/dev/null | tail -1 | cut -c1-200; done
python -m pytest tests/test_gpu_extract.py -q -m gpu -k "throughput or batch_equals" 2>&1 | tail -2
ORBX_DEVICE_CHUNK=32 python -m pytest tests/test_gpu_extract.py tests/test_gpu_gather.py -q -m gpu 2>&1 | tail -2
