/* A plain C99 consumer of include/chanamq_gpu.h, built and run by tests/test_abi_cpu.py on a machine without a GPU:
 * the host half of the boundary (bind / unbind / commit / queue numbering / status codes) through the same symbols a
 * cgo or JNA binding would use.  A host-only table (device -1) compiles and reports; routing on it must fail
 * loudly — there is no CPU matcher. */
#include <stdio.h>
#include <string.h>

#include "chanamq_gpu.h"

#define CHECK(x) do { if (!(x)) { printf("FAILED line %d: %s (%s)\n", __LINE__, #x, chq_last_error()); return 1; } } while (0)

static int32_t bind(chq_table* t, const char* key, const char* queue, uint8_t durable) {
  return chq_bind(t, (const uint8_t*)key, (uint32_t)strlen(key), (const uint8_t*)queue, (uint32_t)strlen(queue), durable);
}

int main(void) {
  chq_table* t = NULL;
  uint32_t n = 0, len = 0, i;
  const uint8_t* name = NULL;
  const char* want[3] = {"alpha", "beta", "gamma"};
  const uint8_t keys[] = "a.b";
  const uint32_t off[2] = {0, 3};
  const uint8_t packed[6] = {0x56, 0x34, 0x12, 0xFF, 0xFF, 0xFF};
  uint32_t ids[2] = {0, 0};
  chq_result r;
  char json[1024];

  CHECK(chq_table_create(&t, CHQ_TOPIC, CHQ_MODE_COMPAT, -1) == CHQ_OK);
  CHECK(bind(t, "a.*", "gamma", 0) == CHQ_OK);
  CHECK(bind(t, "a.b", "alpha", 1) == CHQ_OK);
  CHECK(bind(t, "#.x", "beta", 0) == CHQ_OK);
  CHECK(bind(t, "a.b", "alpha", 1) == CHQ_OK);             /* the matcher holds a set: QueueMatcher.scala:443-446 */
  CHECK(bind(t, ".", "zz", 0) == CHQ_E_ZERO_WORDS);        /* the reference throws here: QueueMatcher.scala:420 */
  CHECK(chq_commit(t) == CHQ_OK);
  CHECK(chq_n_queues(t, &n) == CHQ_OK && n == 3);
  for (i = 0; i < 3; ++i) {                                /* ids are name ranks */
    CHECK(chq_queue_name(t, i, &name, &len) == CHQ_OK);
    CHECK(len == strlen(want[i]) && memcmp(name, want[i], len) == 0);
  }
  CHECK(chq_queue_name(t, 3, &name, &len) == CHQ_E_ARG);
  CHECK(chq_route(t, keys, off, 1, NULL, NULL, &r) == CHQ_E_CUDA);   /* no CPU path */
  CHECK(strlen(chq_last_error()) > 0);
  CHECK(chq_stats(t, json, sizeof json) == CHQ_OK && strstr(json, "\"bindings\": 3") != NULL);
  CHECK(chq_set_ids_format(t, CHQ_IDS_U24) == CHQ_OK && chq_set_ids_format(t, 7) == CHQ_E_ARG);
  chq_unpack_u24(packed, 2, ids);
  CHECK(ids[0] == 0x123456u && ids[1] == 0xFFFFFFu);
  CHECK(chq_table_destroy(t) == CHQ_OK);
  printf("ok\n");
  return 0;
}
