#!/bin/bash
export PYTHONPATH=.
timeout 120 python bench/prof_elem.py > gpurun_out/prof_elem_plain.log 2>&1 || { tail -5 gpurun_out/prof_elem_plain.log; exit 1; }
timeout 400 ncu --set full --clock-control none --import-source on -k regex:"ln_fwd_vec|ln_bwd_vec|segbn_sums|segbn_apply|tc_tile_kernel" --launch-skip 6 -c 14 -o gpurun_out/prof_elem2 -f python bench/prof_elem.py > gpurun_out/prof_elem_ncu.log 2>&1
tail -3 gpurun_out/prof_elem_ncu.log; ls -la gpurun_out/prof_elem2.ncu-rep
