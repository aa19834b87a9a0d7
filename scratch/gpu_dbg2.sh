echo "--- default"; python scratch/dbg_case2.py 1 77 2>&1 | sed -n 2,3p
echo "--- PDL=0"; MDD_PDL=0 python scratch/dbg_case2.py 1 77 2>&1 | sed -n 2,3p
echo "--- SPLIT=0"; MDD_CTC_SPLIT=0 python scratch/dbg_case2.py 1 77 2>&1 | sed -n 2,3p
echo "--- TEAM/DUAL off"; MDD_CTC_DUAL_MAX_PAIRS=0 MDD_CTC_TEAM_MAX_PAIRS=0 python scratch/dbg_case2.py 1 77 2>&1 | sed -n 2,3p
echo "--- DUAL forced up to 592"; MDD_CTC_DUAL_MAX_PAIRS=592 python scratch/dbg_case2.py 1 77 2>&1 | sed -n 2,3p
