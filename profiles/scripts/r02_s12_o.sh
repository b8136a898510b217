# session 12, call O: product build against the measurement build of the same source on the dense-rule probe (same box)
for rep in 1 2; do
echo "== product" | tee -a gpurun_out/dense_rule2.txt
python profiles/scripts/dense_rule_probe.py 2>&1 | head -8 | tee -a gpurun_out/dense_rule2.txt
echo "== measurement build, defaults" | tee -a gpurun_out/dense_rule2.txt
DH_B200_LIB=profiles/trace_build/libdh_tune.so python profiles/scripts/dense_rule_probe.py 2>&1 | head -8 | tee -a gpurun_out/dense_rule2.txt
done
