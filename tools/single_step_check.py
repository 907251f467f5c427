import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import parity
from gym_eldorado_b200 import capi, layout
from oracle import port
L = capi.load(os.path.abspath(sys.argv[1]))
d = parity.run_parity(L, port.OracleVec, 256, 777, layout.HARD, 20, 600, "M")
d += parity.run_parity(L, port.OracleVec, 256, 12345, layout.EASY, 100000, 800, "A")
d += parity.run_parity(L, port.OracleVec, 128, 99, layout.HARD, 8, 400, "M", n_players=2)
print("single-step parity ok, dones", d)
