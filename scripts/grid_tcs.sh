for shape in 42,4096,14336 42,4096,4096 46,3200,8640 42,2048,8192 42,14336,4096; do
 for s in 1 2 4 8; do for r in 128 112 100 96 88 80; do
  TAG="S=$s R=$r" VLUT_TCS_SPLIT=$s VLUT_TCS_ROWS=$r timeout 60 python scripts/time_tcs.py $shape 2>&1 | tail -1
 done; done
done
