timeout 900 python -m pytest tests/test_env_properties_gpu.py -x -q -m gpu -k "step_host" 2>&1 | tail -5
python profiles/tools/e2e_legs.py 2>&1 | tail -1
