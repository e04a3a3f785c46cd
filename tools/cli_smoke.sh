set -e
cd $GRAFT_REPO_ROOT
mkdir -p /tmp/run && cd /tmp/run
R=$GRAFT_REPO_ROOT
echo "== 1024x1024 full"
( time LBM_REPORT=1 $R/d2q9-bgk $R/inputs/input_1024x1024.params $R/inputs/obstacles_1024x1024.dat ) 2>&1 | tail -14
python $R/tools/check.py --help > /dev/null 2>&1 || true
ls -la final_state.dat av_vels.dat
sed 's/^20000$/2000/' $R/inputs/input_16384x1024.params > p16k.params
echo "== 16384x1024 2000 iters 1 GPU"
( time LBM_REPORT=1 $R/d2q9-bgk p16k.params $R/inputs/obstacles_16384x1024.dat ) 2>&1 | tail -14
ls -la final_state.dat
md5sum final_state.dat av_vels.dat > /tmp/run/md5_1
NG=$(nvidia-smi -L | wc -l)
echo "== 16384x1024 2000 iters $NG GPUs"
( time LBM_NGPUS=all LBM_REPORT=1 $R/d2q9-bgk p16k.params $R/inputs/obstacles_16384x1024.dat ) 2>&1 | tail -14
md5sum final_state.dat av_vels.dat; cat /tmp/run/md5_1
