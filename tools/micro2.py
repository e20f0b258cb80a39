import sys; sys.path.insert(0,'/root/repo/tools'); sys.path.insert(0,'/root/repo')
from sweep_micro import run
for args in [(32768, 96), (8192, 8192), (2048, 2048)]:
    for nw in (1, 2, 4, 8):
        run(*args, nw=nw)
run(32768, 96, cap=10, nw=4)
run(32768, 96, cap=5, nw=4)
run(8192, 8192, cap=10, nw=4)
