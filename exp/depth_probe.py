import sys; sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import __graft_entry__ as g
rt=g.load_package()
import test_gpu_parity as t
for pl,lv in ((32768,3),(262144,2),(1048576,1)):
    v,i=t.deep_tree_mesh(rt,pl,lv)
    try:
        sc=rt.Scene(rt.SceneInputs([(v,i)],[(0,rt.TYPE_DIFFUSE,rt.IDENTITY34)],None),device_id=0)
        print(pl,lv,"tris",i.shape[0],"depth",sc.stats()["max_tree_depth"]); sc.close()
    except rt.RtError as e:
        print(pl,lv,"tris",i.shape[0],"REFUSED",e)
