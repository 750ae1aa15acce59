# A/B of tools/ab/*.so on the full C4 frame and on rank 0's share of an 8-GPU frame
cd $GRAFT_REPO_ROOT
echo "== full frame"; bash tools/gpu_ab.sh c4 10 0 | grep " ms "
echo "== 1/8 share";  bash tools/gpu_ab.sh c4 10 8 | grep " ms "
