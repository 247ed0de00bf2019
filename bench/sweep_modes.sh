export DRN_NRES=2
for shape in "14 256 1024 1" "7 512 2048 1" "28 128 512 1" "14 256 256 3"; do set -- $shape
for ms in "0 0" "0 1" "1 0" "1 1" "2 1"; do set -- $shape $ms
python bench/prof_one.py fprop 128 $1 $2 $3 $4 1 $5 $6 5 | cut -c1-100
done; done
