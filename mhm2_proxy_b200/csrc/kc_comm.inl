// Multi-shard supermer exchange (stub until the NCCL path lands). Included at the end of kc_api.cu.
int kc_exchange(kc_handle *h) { return fail(h, KC_ERR_COMM, "multi-shard exchange not initialised: call kc_comm_init"); }
void kc_comm_destroy(kc_handle *) {}
extern "C" int kc_comm_unique_id(uint8_t id128[128]) { (void)id128; return KC_ERR_COMM; }
extern "C" int kc_comm_init(kc_handle *h, const uint8_t id128[128]) { (void)id128; return fail(h, KC_ERR_COMM, "NCCL exchange not built yet"); }
