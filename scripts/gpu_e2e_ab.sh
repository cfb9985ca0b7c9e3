for rep in 1 2; do
for v in base.so stage.so; do
  echo "== $v (default threads)"; FLIVO_LIB=variants/$v python scripts/dbg_e2e.py 2>&1 | tail -2
  echo "== $v (16 threads)"; FLIVO_HOST_THREADS=16 FLIVO_LIB=variants/$v python scripts/dbg_e2e.py 2>&1 | tail -2
done; done
