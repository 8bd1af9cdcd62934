cd $GRAFT_REPO_ROOT
for g in 0 1; do echo -n "small_grid=$g: "; IVX_FINALIZE_SMALL_GRID=$g timeout 200 python tools/overlap_probe.py 2>&1 | tail -3; done
