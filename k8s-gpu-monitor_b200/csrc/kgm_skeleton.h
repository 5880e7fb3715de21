// Constant chunks ("skeleton") of the two line templates, rendered once per node / per
// container when the fleet is loaded (gpu.py:136-143 and gpu.py:206-216 == 240-250).
// Pure host C++: used by kgm_api.cu and by the CPU unit tests.
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "kgm_common.cuh"
#include "kgm_fast.cuh"
#include "kgm_poll2_leaf.cuh"

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

// ---- template block of one node, for the fused kernel's fast path (kgm_fast.cuh) -------------
// header | container directory | node pieces | per gated container its 5 variable pieces
struct TbContIn {
  bool gated;
  std::string cid, ns, pod, user;
  uint32_t vis_begin, nvis;
  uint32_t fixed;             // SkelPool::pod_pieces() of the same strings
};

// Appends the block to `out` (16-byte aligned start) and returns true, or leaves `out`
// untouched and returns false if a length does not fit the block's 16-bit fields.
inline bool build_template_block(const std::string &host, uint32_t node_fixed,
                                 const std::vector<TbContIn> &conts, std::vector<uint8_t> &out,
                                 size_t *start_out, size_t *len_out) {
  const uint32_t ncn = (uint32_t)conts.size();
  const std::string n0 = "node_gpu,host=" + host + ",gpu_id=", n5 = ",node=\"" + host + "\",gid=";
  fast::TbHeader hd{};
  std::vector<fast::TbCont> dir(ncn);
  std::vector<uint8_t> pieces;
  const size_t dir_end = sizeof(fast::TbHeader) + sizeof(fast::TbCont) * (size_t)ncn;
  bool ok = node_fixed <= 60000 && dir_end + n0.size() + n5.size() <= 60000;
  hd.nc = ncn;
  hd.n0_off = (uint16_t)dir_end; hd.n0_len = (uint16_t)n0.size();
  pieces.insert(pieces.end(), n0.begin(), n0.end());
  hd.n5_off = (uint16_t)(dir_end + pieces.size()); hd.n5_len = (uint16_t)n5.size();
  pieces.insert(pieces.end(), n5.begin(), n5.end());
  hd.fixed = (uint16_t)node_fixed;
  for (uint32_t c = 0; c < ncn && ok; ++c) {
    const TbContIn &in = conts[c];
    fast::TbCont &t = dir[c];
    t = fast::TbCont{};
    t.gated = in.gated ? 1 : 0;
    t.vis_begin = in.vis_begin;
    t.nvis = in.nvis;
    t.blk_off = (uint32_t)(dir_end + pieces.size());
    if (!in.gated) continue;
    const std::string pc[5] = {"pod_gpu,host=" + host + ",namespace=" + in.ns + ",gpu_id=",
                               ",containerid=" + in.cid + ",pid_t=",
                               ",user_name_t=" + in.user + " procressname=\"",
                               "\",podname=\"" + in.pod + "\",pid=\"",
                               "\",node=\"" + host + "\",user_name=\"" + in.user + "\",gpu_util="};
    for (int i = 0; i < 5; ++i) {
      if (pc[i].size() > 60000) ok = false;
      t.plen[i] = (uint16_t)pc[i].size();
      pieces.insert(pieces.end(), pc[i].begin(), pc[i].end());
    }
    if (in.fixed > 60000) ok = false;
    t.fixed = (uint16_t)in.fixed;
  }
  const size_t total = dir_end + pieces.size();
  if (!ok || total > 96 * 1024) return false;
  while (out.size() & 15) out.push_back(0);
  *start_out = out.size();
  *len_out = total;
  const uint8_t *hp = reinterpret_cast<const uint8_t *>(&hd);
  out.insert(out.end(), hp, hp + sizeof(hd));
  const uint8_t *dp = reinterpret_cast<const uint8_t *>(dir.data());
  out.insert(out.end(), dp, dp + sizeof(fast::TbCont) * (size_t)ncn);
  out.insert(out.end(), pieces.begin(), pieces.end());
  return true;
}

// ---- template slots of k_poll2 (kgm_poll2_leaf.cuh) -------------------------------------------
// One slot per node (the two pieces of its node_gpu lines) and per gated container (the five
// pieces of its pod_gpu lines): header | piece bytes, every piece on a 16-byte boundary.
//   header  8 run words relative to the slot start (piece i), ..., u16 constant bytes of a whole
//           line incl. its '\n' at byte 46
// Appends the slot to `pool` and returns true; false (pool untouched) if a piece is too long.
inline bool build_slot(bool pod, const std::string *pieces, int npieces, std::vector<uint8_t> &pool,
                       size_t *start_out, size_t *len_out) {
  using namespace p2;
  uint32_t W[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  uint32_t at = SLOT_HEADER;
  for (int i = 0; i < npieces; ++i) {
    if (pieces[i].size() > 2000) return false;
    W[i] = rw(at, (uint32_t)pieces[i].size());
    at += ((uint32_t)pieces[i].size() + 15u) & ~15u;
  }
  if (at > 8192) return false;
  uint16_t tail[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  uint32_t cst = 0;
  const int ns = pod ? (int)POD_SEGS : (int)NODE_SEGS;
  for (int k = 0; k < ns; ++k) {
    const uint32_t e = pod ? seg_pod(k) : seg_node(k);
    if ((e >> 6) == 0u) cst += rw_len(W[e & 63u]);
    else if ((e >> 6) == 1u) cst += k_len(e & 63u);
  }
  tail[7] = (uint16_t)cst;
  while (pool.size() & 15) pool.push_back(0);
  *start_out = pool.size();
  const uint8_t *wp = reinterpret_cast<const uint8_t *>(W), *tp = reinterpret_cast<const uint8_t *>(tail);
  pool.insert(pool.end(), wp, wp + 32);
  pool.insert(pool.end(), tp, tp + 16);
  for (int i = 0; i < npieces; ++i) {
    pool.insert(pool.end(), pieces[i].begin(), pieces[i].end());
    while (pool.size() & 15) pool.push_back(0);
  }
  *len_out = pool.size() - *start_out;
  return true;
}

inline bool build_node_slot(const std::string &host, std::vector<uint8_t> &pool, size_t *start, size_t *len) {
  const std::string pc[2] = {"node_gpu,host=" + host + ",gpu_id=", ",node=\"" + host + "\",gid="};
  return build_slot(false, pc, 2, pool, start, len);
}

inline bool build_pod_slot(const std::string &host, const std::string &ns, const std::string &cid,
                           const std::string &user, const std::string &pod, std::vector<uint8_t> &pool,
                           size_t *start, size_t *len) {
  const std::string pc[5] = {"pod_gpu,host=" + host + ",namespace=" + ns + ",gpu_id=",
                             ",containerid=" + cid + ",pid_t=",
                             ",user_name_t=" + user + " procressname=\"",
                             "\",podname=\"" + pod + "\",pid=\"",
                             "\",node=\"" + host + "\",user_name=\"" + user + "\",gpu_util="};
  return build_slot(true, pc, 5, pool, start, len);
}

}  // namespace kgm
