# session 12, call M: breakeven of the per-unit tile renderer against the staged paths (dense_rule_probe.py)
export DH_B200_LIB=profiles/trace_build/libdh_tune.so
echo "== always tiles (DH_G_DENSE_MIN=1 DH_G_DENSE_COST16=0)" | tee -a gpurun_out/dense_rule.txt
DH_G_DENSE_MIN=1 DH_G_DENSE_COST16=0 python profiles/scripts/dense_rule_probe.py 2>&1 | tee -a gpurun_out/dense_rule.txt
echo "== never tiles (DH_G_DENSE_COST16=2000000000)" | tee -a gpurun_out/dense_rule.txt
DH_G_DENSE_COST16=2000000000 python profiles/scripts/dense_rule_probe.py 2>&1 | tee -a gpurun_out/dense_rule.txt
echo "== rule as built (min 8, cost16 32)" | tee -a gpurun_out/dense_rule.txt
python profiles/scripts/dense_rule_probe.py 2>&1 | tee -a gpurun_out/dense_rule.txt
