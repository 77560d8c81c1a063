// placeholder, replaced below
#include "heat_ctx.cuh"
namespace heat {
int comm_unique_id(uint8_t *) { return fail("comm not built yet"); }
int comm_init(Ctx *, int, int, const uint8_t *) { return fail("comm not built yet"); }
int comm_sync_items(Ctx *) { return fail("comm not built yet"); }
void comm_destroy(Ctx *) {}
}
