for k in 5 15 24; do for tr in 64 128 256 384 512; do
  echo -n "k=$k TR=$tr "; RAILS_LS_TILE_ROWS=$tr python scripts/chain_bench.py --grid 256 --k $k 2>&1 | tail -1 | cut -c1-120
done; done
