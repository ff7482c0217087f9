"""mnist_async_sharding_greedy worker rank.  usage (via run.sh): python3 worker.py -np <num_workers>"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ddl_b200.entry import worker_main

if __name__ == "__main__":
    worker_main("mnist_async_sharding_greedy")
