import sys; sys.path.insert(0,'.')
from lgattn_pkg import pkg
for (b,n) in [(64,2048),(16,4096),(8,4096),(10,2048),(27,1024),(128,1024),(48,300),(40,512),(100,100),(80,777),(24,1100),(64,640),(256,512),(30,2000)]:
    print(b,n, pkg._binding.describe_launch(b,4,n,n))
