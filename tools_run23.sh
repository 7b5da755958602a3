python -m pytest tests -m gpu -q -x 2>&1 | tail -3
python tools/kbench.py single2d 4096
python tools/kbench.py ros2d 4096
python tools/kbench.py ros2d 1024
