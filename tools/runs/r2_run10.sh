set -x
mkdir -p gpurun_out
for opt in "pdl=1" "rrdb_ring=1" "pdl=0" "rrdb_ring=1 pdl=0"; do
  timeout 900 python tools/rrdb_bench.py 23 $opt 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$opt', 'tile %.2f ms' % d['tile_532']['ms'], 'tiled %.1f ms' % d['frame_1080p_tiled']['ms'], 'whole %.1f ms' % d['frame_1080p_whole']['ms'])"
done
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw --format=csv
