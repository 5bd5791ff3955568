import numpy as np
p=np.load('gpurun_out/mega_profile.npy'); us=1/1965.
q=p[:,:,1,:]
last=(q[...,4]>q[...,6])
w=(q[...,4]-q[...,6])[last]*us; l=(q[...,3]-q[...,4])[last]*us
print("last CTAs:",last.sum(),"weights mean %.2f max %.2f; loop mean %.2f max %.2f"%(w.mean(),w.max(),l.mean(),l.max()))
# 2-head vs 1-head
