for ns in 0 20 64 200; do echo "== DRN_SPIN_NS=$ns"; export DRN_SPIN_NS=$ns
python bench/prof_pre.py 2>&1 | head -4
python bench/tune_igemm.py 2>&1 | sed "s/.*-> //" | awk '{print $NF-1+1, $(NF-1)}' | tr '\n' ' ' ; echo
done
