for B in 512 256; do
for w in 7 8; do for n in 3 4 5 6 8; do
echo -n "B=$B WPS=$w NSEG=$n: "; MDD_CTC_PAIR_WPS=$w MDD_CTC_DUAL_NSEG=$n python scratch/time_one.py 1000 $B 42 120 | tail -1 | cut -c1-50
done; done; done
