import sys, numpy as np
sys.path.insert(0, '/root/repo')
from acclimate_b200 import engine, synth
net = synth.make_config("eora_flood")
e = engine.Engine(net); sc = engine.Scenario(net)
e.profile(True)
for t in range(22):
    sc.apply(e, t); e.step(); e.sync()
    kt = e.kernel_times(reset=True)
    print(t, "sorted sellers", int(e.read_state("firm_sorted_sums").sum()), "seller_production ms", round(kt["seller_production"][0], 3), flush=True)
e.close()
