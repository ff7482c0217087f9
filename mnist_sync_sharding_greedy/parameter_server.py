"""mnist_sync_sharding_greedy parameter server rank.  usage (via run.sh): python3 parameter_server.py -np <num_ps>"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ddl_b200.entry import ps_main

if __name__ == "__main__":
    ps_main("mnist_sync_sharding_greedy")
