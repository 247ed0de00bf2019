# per-shape wgrad timings with the default split policy and forced waves, then the whole step
for w in 0 1 2; do export DRN_WGRAD_WAVES=$w; echo "== waves $w (0 = policy)"
for shape in "56 64 64 3" "56 64 256 1" "56 256 64 1" "28 128 128 3" "28 512 128 1" "14 256 256 3" "14 1024 256 1" "14 256 1024 1" "7 512 512 3" "7 2048 512 1" "7 512 2048 1"; do set -- $shape
python bench/prof_one.py wgrad 128 $1 $2 $3 $4 1 0 0 5 | cut -c1-95; done; done
unset DRN_WGRAD_WAVES
timeout 200 python bench.py --steps 30 --warmup 5 2>/dev/null | tail -1 | cut -c1-120
timeout 200 python bench.py --steps 30 --warmup 5 2>/dev/null | tail -1 | cut -c1-120
