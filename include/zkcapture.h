/*
 * zkcapture.h — on-disk container for captured ZooKeeper connections ("zkcap", version 1).
 *
 * The reference has no capture format (it only ever sees live sockets); this one is ours.  It
 * holds exactly what zkb_replay_host (zkbatch.h) takes: the client->server and server->client
 * byte streams of many connections, cut into *segments*, with one row of metadata per segment.
 * The file is meant to be mmap'ed: every section starts on a 4096-byte boundary, so the two
 * byte sections can be pinned (cudaHostRegister) and handed to the codec as they lie.
 *
 * A segment is a stretch of one connection: bytes [req_off[i], req_off[i+1]) of the request
 * section and [rsp_off[i], rsp_off[i+1]) of the response section.  Both cuts fall on frame
 * boundaries.  The codec treats every segment as a session of its own, which is how the
 * container expresses what a flat pair of streams cannot: the order between the two
 * directions.  The reference pairs a reply with the opcode stored under its xid *at the time
 * the reply arrives* (ZKEncodeStream stores, lib/zk-streams.js:145; readResponse looks up,
 * lib/zk-buffer.js:288-290), so a connection that re-uses xids decodes correctly only if
 * replies are matched with the requests sent before them.  A capture tool therefore cuts a
 * connection wherever it was quiescent (every request answered) before xids repeat: inside a
 * segment xids are unique, and "last write wins" over the segment's requests is exact.  The
 * real client never re-uses an xid on a connection (lib/connection-fsm.js:74-76 counts up), so
 * its captures need one segment per connection.
 *
 * All integers little-endian.  No function is declared here: reader and writer are a few
 * lines in any language (node-zkstream_b200/capture.py is the one the tests use).
 */
#ifndef ZKCAPTURE_H
#define ZKCAPTURE_H

#include <stdint.h>

#define ZKCAP_MAGIC    "ZKCAP\0\0\1"   /* 8 bytes; the last one is the format version */
#define ZKCAP_VERSION  1u
#define ZKCAP_ALIGN    4096u

/* Byte offsets are from the start of the file; a section with offset 0 is absent. */
typedef struct zkcap_header {
	char     magic[8];
	uint32_t version;         /* ZKCAP_VERSION */
	uint32_t nsegments;
	uint64_t req_bytes;       /* length of the request byte section */
	uint64_t rsp_bytes;       /* length of the response byte section */
	uint64_t off_req_off;     /* uint64[nsegments + 1]  segment starts in the request section  (zkb_sessions.off) */
	uint64_t off_rsp_off;     /* uint64[nsegments + 1]  segment starts in the response section (zkb_sessions.off) */
	uint64_t off_conn;        /* uint32[nsegments]      connection the segment belongs to; equal ids are adjacent and in stream order */
	uint64_t off_xid_base;    /* int32[nsegments]       first xid tracked for the segment (zkb_replay_host xid_base) */
	uint64_t off_xid_count;   /* uint32[nsegments]      number of consecutive xids tracked   (zkb_replay_host xid_count) */
	uint64_t off_req_hs;      /* uint32[nsegments]      request frames read in 'handshaking' state: 1 for a connection's first segment, else 0 */
	uint64_t off_rsp_hs;      /* uint32[nsegments]      same for the response direction */
	uint64_t off_req;         /* uint8[req_bytes] */
	uint64_t off_rsp;         /* uint8[rsp_bytes] */
	uint64_t file_bytes;      /* total length, for truncation checks */
	uint8_t  reserved[16];
} zkcap_header;               /* 128 bytes */

#endif
