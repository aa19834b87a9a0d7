python scratch/ncu_others.py > gpurun_out/ncu_others_plain.log 2>&1 && ncu --set full --clock-control none -k 'regex:(beam_search|edit_align|argmax_rows|collapse|prefix_score|prefix_select|prefix_init|corpus_counts|row_lse|ctc_exact|ctc_team|exp_transpose)' -c 60 -f -o /tmp/prof_others_r2 python scratch/ncu_others.py > gpurun_out/ncu_others_run.log 2>&1
ncu -i /tmp/prof_others_r2.ncu-rep --page raw --csv > gpurun_out/prof_others_r2_raw.csv 2>/dev/null
ls -la /tmp/prof_others_r2.ncu-rep gpurun_out/prof_others_r2_raw.csv
