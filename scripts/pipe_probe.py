import sys, json; sys.path.insert(0, __import__('os').path.dirname(__import__('os').path.dirname(__import__('os').path.abspath(__file__))))
import torch, bench
print(json.dumps(bench._pipeline_wall(torch.device("cuda:0")), indent=1))
print(json.dumps(bench._pipeline_wall(torch.device("cuda:0")), indent=1))
