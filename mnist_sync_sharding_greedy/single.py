"""Single-process trainer (local optimizer, no parameter server).  usage: python3 single.py"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from ddl_b200.entry import single_main

if __name__ == "__main__":
    single_main("mnist_sync_sharding_greedy")
