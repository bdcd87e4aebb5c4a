cd $GRAFT_REPO_ROOT
timeout 75 python -m pytest tests/test_observers.py tests/test_temperature.py -m gpu -s -q -k "test_device_processors_match_oracle or test_rk_steps_with_temperature" 2>&1 | grep -E "tolerance\]|passed|failed" | head -40
