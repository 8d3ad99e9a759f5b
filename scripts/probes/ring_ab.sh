for a in "4000000 128 100 3" "4000000 128 72 3" "4000000 128 256 3" "4000000 96 100 3" "4000000 112 200 3" "4000000 96 600 2" "2000000 128 1000 2 predict" "2000000 128 1000 2" "4000000 128 360 2"; do
  echo "old: $(FMLK_LIB_PATH=$PWD/scripts/probes/lib_old.so timeout 120 python scripts/run_assign.py $a 2>&1 | tail -1 | cut -c1-90)"
  echo "new: $(timeout 120 python scripts/run_assign.py $a 2>&1 | tail -1 | sed 's/regs=.*\(ring=[0-9]\).*/\1/' | cut -c1-110)"
done
