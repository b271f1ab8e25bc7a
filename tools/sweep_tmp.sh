for shp in "64 20 20 768 768 1 1" "64 40 40 384 384 1 1" "64 80 80 192 192 1 1" "64 40 40 192 192 1 1"; do
 for tn in 64 96 128 192 256; do for mb in 1 2; do
  timeout 60 python tools/one_op.py $shp --tile-n $tn --mblk $mb 2>&1 | tail -1 | cut -c1-230
 done; done
done
for shp in "64 40 40 192 192 3 1" "64 80 80 96 96 3 1" "64 20 20 384 384 3 1"; do
 for tn in 64 96 128 192; do for mb in 1 2 4; do
  timeout 60 python tools/one_op.py $shp --tile-n $tn --mblk $mb 2>&1 | tail -1 | cut -c1-230
 done; done
done
