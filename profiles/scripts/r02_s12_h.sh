# session 12, call H: what the T >= 256 selector of the tile renderer does to SPARSE data in widely padded slot arrays
echo "== product (tile renderer from T >= 256)" | tee -a gpurun_out/padded_t.txt
python profiles/scripts/padded_t_probe.py 2>&1 | tee -a gpurun_out/padded_t.txt
echo "== measurement build, DH_G_TILES=0 (staged renderer, lean layout from T >= 256)" | tee -a gpurun_out/padded_t.txt
DH_B200_LIB=profiles/trace_build/libdh_tune.so DH_G_TILES=0 python profiles/scripts/padded_t_probe.py 2>&1 | tee -a gpurun_out/padded_t.txt
echo "== measurement build, DH_G_TILES=0 DH_G_LEAN=0" | tee -a gpurun_out/padded_t.txt
DH_B200_LIB=profiles/trace_build/libdh_tune.so DH_G_TILES=0 DH_G_LEAN=0 python profiles/scripts/padded_t_probe.py 2>&1 | tee -a gpurun_out/padded_t.txt
