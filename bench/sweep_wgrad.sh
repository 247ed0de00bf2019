for w in 1 2 3; do export DRN_WGRAD_WAVES=$w; echo "== waves $w"
for shape in "56 64 64 3" "56 64 256 1" "56 256 64 1" "28 128 128 3" "28 512 128 1" "14 256 256 3" "14 1024 256 1" "14 256 1024 1" "7 512 512 3" "7 2048 512 1" "7 512 2048 1"; do set -- $shape
python bench/prof_one.py wgrad 128 $1 $2 $3 $4 1 0 0 5 | cut -c1-95; done; done
timeout 200 python bench.py --steps 10 --warmup 3 2>&1 | tail -1 | cut -c1-200
DRN_WGRAD_WAVES=1 timeout 200 python bench.py --steps 10 --warmup 3 2>&1 | tail -1 | cut -c1-200
