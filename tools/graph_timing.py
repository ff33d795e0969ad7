import sys, time, numpy as np
sys.path.insert(0,'/root/repo')
import feelpp_b200 as fpp
eng=fpp.Engine(0)
mesh=fpp.unitCube(110,0.15,False)
t0=time.time(); Xh=fpp.Pch(mesh,2); t1=time.time()
ia,ja=Xh.graph(); t2=time.time()
eng.space(Xh); eng.sync(); t3=time.time()
dia,dja=eng.graph(Xh); t4=time.time()
print("doftable %.2fs host graph %.2fs | device graph incl. D2H of ia/ja %.2fs | equal: %s %s nnz %d"%(t1-t0,t2-t1,t4-t3,np.array_equal(ia,dia),np.array_equal(ja,dja),len(ja)))
t5=time.time(); dia,dja=eng.graph(Xh); t6=time.time(); print("second call %.2fs"%(t6-t5))
