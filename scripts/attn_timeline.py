import sys,statistics
names="kIss vIss pfullA pfullB issA issB compA compB sSeenA sSeenB pArrA pArrB vfull kfullN".split()
for f in sys.argv[1:]:
    rows=[[int(x) for x in l.split()] for l in open(f) if not l.startswith('#')]
    rows=[r for r in rows if all(x>0 for x in r[1:])][10:150]
    col=lambda n:[r[1+names.index(n)] for r in rows]
    per=statistics.median([b-a for a,b in zip(col('pArrA'),col('pArrA')[1:])])
    def d(a,b,shift=0):
        A=col(a);B=col(b)
        if shift: return statistics.median([y-x for x,y in zip(A[:-shift],B[shift:])])
        return statistics.median([y-x for x,y in zip(A,B)])
    print(f.split('/')[-1],'period',per)
    print('  softA: compat done->S seen %d | S seen->P arrive %d | P arrive->next compat done %d'%(d('compA','sSeenA'),d('sSeenA','pArrA'),d('pArrA','compA',1)))
    print('  softB: compat done->S seen %d | S seen->P arrive %d | P arrive->next compat done %d'%(d('compB','sSeenB'),d('sSeenB','pArrB'),d('pArrB','compB',1)))
    print('  mma: P_A arrive->seen %d | seen A->issued A %d | issued A->P_B seen %d | P_B arr->seen %d | seen B->issued B %d | issued B->next P_A seen %d'%(d('pArrA','pfullA'),d('pfullA','issA'),d('issA','pfullB'),d('pArrB','pfullB'),d('pfullB','issB'),d('issB','pfullA',1)))
    print('  S_A latency: issued A(t) -> S seen A(t+1) %d ; S_B: issued B(t) -> S seen B(t+1) %d'%(d('issA','sSeenA',1),d('issB','sSeenB',1)))
    print('  mma detail: pfullA->vfull %d | vfull->kfullN (PV_A issue) %d | kfullN->issA (QK_A issue) %d'%(d('pfullA','vfull'),d('vfull','kfullN'),d('kfullN','issA')))
