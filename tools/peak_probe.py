import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200ba
pb = b200ba.synthetic.make_problem(8, 300, 1200, seed=1, window=8)
eng = b200ba.Engine(0)
eng.set_problem(*pb.args())
for bps in (8, 4, 2, 1):
    os.environ["B200BA_PEAK_BLOCKS_PER_SM"] = str(bps)
    print("blocks/SM", bps, "warps/SM", bps * 8, "TFLOP/s %.2f" % eng.fp64_peak_tflops())
