// Constant chunks ("skeleton") of the two line templates, rendered once per node / per
// container when the fleet is loaded (gpu.py:136-143 and gpu.py:206-216 == 240-250).
// Pure host C++: used by kgm_api.cu and by the CPU unit tests.
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "kgm_common.cuh"

namespace kgm {

struct SkelPool {
  std::vector<uint8_t> bytes;
  Piece n1, n2, n3, n4, n6, p2, p5, p7;      // chunks that do not depend on any name
  SkelPool() {
    n1 = add(" gpu_util="); n2 = add(",mem_util="); n3 = add(",mem_used=");
    n4 = add(",mem_total="); n6 = add(",mem_free=");
    p2 = add(",gpu_util_t="); p5 = add("\",gpu_id_v=\""); p7 = add(",used_memory=");
  }
  Piece add(const std::string &s) {           // every chunk starts 16-byte aligned
    while (bytes.size() & 15) bytes.push_back(0);
    Piece p{(uint32_t)bytes.size(), (uint32_t)s.size()};
    bytes.insert(bytes.end(), s.begin(), s.end());
    return p;
  }
  // node_gpu,host=H,gpu_id=<K> gpu_util=<U>,mem_util=<F>,mem_used=<MU>,mem_total=<MT>,node="H",gid=<K>,mem_free=<MF>
  uint32_t node_pieces(const std::string &host, Piece *out) {
    out[0] = add("node_gpu,host=" + host + ",gpu_id=");
    out[1] = n1; out[2] = n2; out[3] = n3; out[4] = n4;
    out[5] = add(",node=\"" + host + "\",gid=");
    out[6] = n6;
    uint32_t fixed = 0;
    for (int i = 0; i < NODE_PIECES; ++i) fixed += out[i].len;
    return fixed;
  }
  // pod_gpu,host=H,namespace=NS,gpu_id=<G>,containerid=C,pid_t=<P>,gpu_util_t=<U>,user_name_t=UN
  //   procressname="<PN>",podname="POD",pid="<P>",gpu_id_v="<G>",node="H",user_name="UN",gpu_util=<U>,used_memory=<UM>
  uint32_t pod_pieces(const std::string &host, const std::string &ns, const std::string &cid,
                      const std::string &user, const std::string &pod, Piece *out) {
    out[0] = add("pod_gpu,host=" + host + ",namespace=" + ns + ",gpu_id=");
    out[1] = add(",containerid=" + cid + ",pid_t=");
    out[2] = p2;
    out[3] = add(",user_name_t=" + user + " procressname=\"");
    out[4] = add("\",podname=\"" + pod + "\",pid=\"");
    out[5] = p5;
    out[6] = add("\",node=\"" + host + "\",user_name=\"" + user + "\",gpu_util=");
    out[7] = p7;
    uint32_t fixed = 0;
    for (int i = 0; i < POD_PIECES; ++i) fixed += out[i].len;
    return fixed;
  }
};

}  // namespace kgm
