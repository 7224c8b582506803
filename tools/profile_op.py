"""Run a few launches of one operator shape for ncu. Usage: profile_op.py <substring of op name>"""
import os, sys
sys.argv = [sys.argv[0]] + sys.argv[1:]
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import importlib.util
