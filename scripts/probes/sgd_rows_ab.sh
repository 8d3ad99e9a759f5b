# group rows of the SGD kernels (a -DFMLK_TUNING build in scripts/probes/lib_tune.so): us per epoch at three batch sizes
export FMLK_LIB_PATH=$PWD/scripts/probes/lib_tune.so
for r in 32 16 8; do
  for shape in "10000000 100 100000 20" "10000000 100 1000000 20" "1000000 100 1000 20"; do
    echo "rows=$r $shape: $(FMLK_TUNE_SGD_ROWS=$r timeout 100 python scripts/measure_sgd.py $shape 2>&1 | tail -1 | sed 's/.*"us_per_epoch": \([0-9.]*\).*"hbm_frac": \([0-9.]*\).*"parity_on_sample": \([a-z]*\).*/us_per_epoch=\1 hbm_frac=\2 parity=\3/')"
  done
done
