set -x
nvidia-smi --query-gpu=name,memory.total --format=csv
cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_somatic_standard.py -x -q -m gpu 2>&1 | tail -40
python __graft_entry__.py --smoke 2>&1 | tail -5
