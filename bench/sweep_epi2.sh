# marginal cost of the epilogue features on chain-critical 1x1 shapes; side ring depth; N tile
for shape in "56 64 256" "28 128 512" "14 256 1024" "7 512 2048" "14 1024 256" "56 256 64"; do set -- $shape
 for ms in "0 0" "0 1" "1 0" "1 1" "2 1"; do set -- $shape $ms
  for nres in 2 4; do
   [ "$4" = "0" ] && [ "$nres" = "4" ] && continue
   echo -n "NRES=$nres "; DRN_NRES=$nres python bench/prof_one.py fprop 128 $1 $2 $3 1 1 $4 $5 5 | cut -c1-100
  done
 done
done
for bn in 64 128 256; do echo -n "force_bn=$bn "; python bench/prof_one.py fprop 128 14 256 1024 1 1 1 1 5 $bn | cut -c1-100; done
for bn in 64 128 256; do echo -n "force_bn=$bn "; python bench/prof_one.py fprop 128 14 1024 256 1 1 2 1 5 $bn | cut -c1-100; done
