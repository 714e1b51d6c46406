# GPU box: sweep the coherent-run length of the query -> lane map (B200_ICP_COH = log2 of the run) on C2
for c in 0 1 2 3 4 5; do
  echo "== B200_ICP_COH=$c"
  B200_ICP_COH=$c python scripts/phase_times.py 2>&1 | grep -A12 "^iters" | sed -n '14,39p' | grep -E "iters|sum|^ +[0-9]" 
done
