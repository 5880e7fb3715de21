"""CPU: the window oracle on a hand-computed case (it is the checker of the GPU stage-3
test; the reference has no counterpart, so this pins the oracle to arithmetic done by hand)."""
from oracle import window_oracle as wo

GPU = ("h\nGPU-a, 10 %, 0 %, 100 MiB, 40 MiB, 60 MiB, 1.50 W\n"
       "GPU-b, [N/A], 0 %, 100 MiB, 7 MiB, 93 MiB, [N/A]\n")


def _doc(tick, util_a, mem_a, pod_mem):
    lines = ['node_gpu,host=n,gpu_id=0 gpu_util=%s,mem_util=1.0,mem_used=%s,mem_total=100,node="n",gid=0,mem_free=1'
             % (util_a, mem_a),
             'node_gpu,host=n,gpu_id=1 gpu_util=,mem_util=7.0,mem_used=7,mem_total=100,node="n",gid=1,mem_free=93',
             'pod_gpu,host=n,namespace=ns,gpu_id=0,containerid=c1,pid_t=5,gpu_util_t=%s,user_name_t=u '
             'procressname="p",podname="pod-u-0",pid="5",gpu_id_v="0",node="n",user_name="u",gpu_util=%s,'
             'used_memory=%s' % (util_a, util_a, pod_mem),
             'pod_gpu,host=n,namespace=ns,gpu_id=None,containerid=c2,pid_t=-,gpu_util_t=0,user_name_t=u '
             'procressname="-",podname="pod-u-1",pid="-",gpu_id_v="None",node="n",user_name="u",gpu_util=0,'
             'used_memory=0']
    return (0, tick, lines, GPU, {"c1": 3, "c2": 4})


def test_hand_computed_windows():
    docs = [_doc(0, "10", "40", "300"), _doc(3, "30", "20", "99999999999"), _doc(1, "20", "60", "100")]
    w = wo.windows(docs, 2)          # ticks 0,1 -> window 0; tick 3 -> window 1
    assert list(w) == [(0, 0, 0, 0), (0, 0, 0, 1), (0, 0, 1, 0), (0, 0, 1, 1), (1, 3, 0, 0), (1, 3, 0, 1),
                       (1, 4, wo.NONE_ID, 0), (1, 4, wo.NONE_ID, 1)]
    g = w[(0, 0, 0, 0)]
    assert g["count"] == 2
    assert (g["gpu_util"]["min"], g["gpu_util"]["max"], g["gpu_util"]["sum"], g["gpu_util"]["last"]) == (10, 20, 30, 20)
    assert g["gpu_util"]["mean"] == 15.0
    assert (g["mem"]["min"], g["mem"]["max"], g["mem"]["last"]) == (40, 60, 60)
    assert g["power"]["sum"] == 300 and g["power"]["count"] == 2          # "1.50 W" -> 150
    g = w[(0, 0, 1, 0)]                                                  # gpu 1: util and power have no digits
    assert g["count"] == 2 and g["gpu_util"]["count"] == 0 and g["power"]["count"] == 0 and g["mem"]["sum"] == 14
    g = w[(1, 3, 0, 1)]
    assert g["mem"]["last"] == wo.SAT == g["mem"]["max"]                  # saturates at 2^32-2
    g = w[(1, 4, wo.NONE_ID, 0)]                                          # idle lines: zeros, no power
    assert g["count"] == 2 and g["gpu_util"]["sum"] == 0 and g["power"]["count"] == 0
