for n in 2000000 5000000 10000000 14200000 19000000; do for m in 1000 2000000000; do
echo "n $n binmin $m: $(KN_BIN_MIN=$m python profiles/run_case.py --n $n --steps 5 --no-solve 2>&1 | tail -1 | cut -c1-40)"
done; done
