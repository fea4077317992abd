// packet_format.h -- the fork's device wire format, restated from the reference's packed structs;
// every struct is 64 bytes (asserted by the reference at main.c:296-302).  Shared by the packet-level entry
// (seed_api.cu: mm2b200_seed_work) and the driver shim (fpgadrv_shim.cu: fpga_writebuf_submit -> fpga_get_retbuf).
#pragma once
#include <stdint.h>

namespace mm2b200 {

// chaindp_sndhdr_t, fpga_chaindp.h:79-87 -- first 64 bytes of a request and of a response packet
struct __attribute__((__packed__)) PktHead { uint32_t magic, size; uint16_t tid, num; uint8_t type, lat; uint8_t reserve1[50]; };
// collect_task_t, fpga_chaindp.h:46-58 -- one per read of a request, followed by seednum minimizers padded to 64 bytes
struct __attribute__((__packed__)) PktTask { int32_t gap_qry, gap_ref, seednum, qlensum; uint32_t read_id, bid; int16_t n_segs; char b; char reserve1[1]; uint64_t mv_a; char reserve2[28]; };
// collect_result_t, fpga_chaindp.h:60-68 -- one per read of a response, followed by n_a new_seed records and n_minipos
// mini_pos words, each padded to 64 bytes; with err_flag = 1 nothing follows and the host redoes the read (map.c:933-944,970-971)
struct __attribute__((__packed__)) PktResult { uint32_t err_flag, read_id, sub_size, n_a, n_minipos, rep_len; char reserve1[40]; };
static_assert(sizeof(PktHead) == 64 && sizeof(PktTask) == 64 && sizeof(PktResult) == 64, "packet headers are 64 bytes");

#define MM2B200_PKT_TYPE 3                     // chaindp_sndhdr_t.type of a chaining request (map.c:301,491)
inline int64_t pkt_up64(int64_t x) { return (x + 63) & ~(int64_t)63; }   // ADDR_ALIGN(x, 64), fpga_chaindp.h:18

} // namespace mm2b200
