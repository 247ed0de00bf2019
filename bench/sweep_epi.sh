for cfg in "2 256" "2 100000" "4 256" "3 100000"; do set -- $cfg; export DRN_NRES=$1 DRN_DEEPK=$2; echo "== NRES=$1 DEEPK=$2";
python bench/prof_one.py fprop 128 28 128 512 1 1 1 1 5 | cut -c1-90
python bench/prof_one.py fprop 128 14 256 1024 1 1 1 1 5 | cut -c1-90
python bench/prof_one.py fprop 128 7 512 2048 1 1 1 1 5 | cut -c1-90
python bench/prof_one.py fprop 128 28 512 128 1 1 2 1 5 | cut -c1-90
python bench/prof_one.py fprop 128 14 1024 256 1 1 2 1 5 | cut -c1-90
python bench/prof_one.py fprop 128 7 2048 512 1 1 2 1 5 | cut -c1-90
python bench/prof_one.py fprop 128 14 256 256 3 1 2 1 5 | cut -c1-90
python bench/prof_one.py fprop 128 7 512 512 3 1 2 1 5 | cut -c1-90
python bench/prof_one.py fprop 128 56 64 256 1 1 1 1 5 | cut -c1-90
done
