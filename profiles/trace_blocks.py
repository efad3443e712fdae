import numpy as np, sys
a=np.fromfile(sys.argv[1],dtype=np.uint64).reshape(-1,3)
t0=a[:,0].astype(np.int64); t1=a[:,1].astype(np.int64); sm=a[:,2].astype(np.int64)
T0=t0.min(); t0-=T0; t1-=T0
dur=t1-t0
print("blocks",len(a),"kernel span us",t1.max()/1e3)
print("block duration us: mean %.1f p50 %.1f p90 %.1f p99 %.1f max %.1f"%(dur.mean()/1e3,np.percentile(dur,50)/1e3,np.percentile(dur,90)/1e3,np.percentile(dur,99)/1e3,dur.max()/1e3))
end_by_sm=np.array([t1[sm==s].max() for s in np.unique(sm)])
print("SMs",len(end_by_sm),"finish time us: min %.1f mean %.1f max %.1f"%(end_by_sm.min()/1e3,end_by_sm.mean()/1e3,end_by_sm.max()/1e3))
busy=np.array([dur[sm==s].sum() for s in np.unique(sm)])
print("sum of block durations per SM (us): min %.0f mean %.0f max %.0f"%(busy.min()/1e3,busy.mean()/1e3,busy.max()/1e3))
# how many blocks start after x% of time
for f in (0.5,0.7,0.8,0.9):
    print("blocks starting after %.0f%% of span: %d ; still running at that time: %d"%(f*100,(t0>f*t1.max()).sum(), ((t0<=f*t1.max())&(t1>f*t1.max())).sum()))
# longest blocks: when do they start
idx=np.argsort(-dur)[:10]
print("longest blocks (id, start us, dur us):",[(int(i),round(t0[i]/1e3,1),round(dur[i]/1e3,1)) for i in idx])
# duration by tile position
nb=len(a); q=np.arange(nb)
for lo in range(0,nb,nb//12): print(lo, "mean dur us %.1f"%(dur[lo:lo+nb//12].mean()/1e3), end=" | ")
print()
