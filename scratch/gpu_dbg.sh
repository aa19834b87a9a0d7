python scratch/dbg_case.py 333 682 62 114
python scratch/dbg_case.py 333 682 42 114
python scratch/dbg_case.py 333 682 62 114 full
MDD_PDL=0 python scratch/dbg_case.py 333 682 62 114
MDD_CTC_SPLIT=0 python scratch/dbg_case.py 333 682 62 114
python scratch/dbg_case.py 333 682 61 114
python scratch/dbg_case.py 333 682 60 114
