for b in 1 2 3 4 6 12 16; do
echo "BATCH $b"
timeout 600 python tools/rrdb_bench.py 23 rrdb_batch=$b 2>/dev/null | tail -1 | python -c "import sys,json; r=json.loads(sys.stdin.read()); print(r['frame_1080p_tiled'])"
done
timeout 600 python -m pytest tests/test_rrdb_gpu.py -x -q -k "batched" 2>&1 | tail -3
