cd $GRAFT_REPO_ROOT
bash tools/ab.sh box4 box5 prev 2>&1 | grep -E "==|value|downsample"
