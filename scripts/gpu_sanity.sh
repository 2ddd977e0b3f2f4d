cd /root/repo
# the between-step mutators with one CUDA-graph launch per step (graphs are dropped when device arrays grow)
B2CUDA_GRAPH=1 timeout 600 python -m pytest tests -m gpu -x -q -k "create or destroy or churn or set_body_type or explode or set_filter" 2>&1 | tail -3
