// cuda_emu.cpp — TEST INFRASTRUCTURE: the fibre scheduler behind cuda_emu.h (see there).
#include "cuda_emu.h"

#include <stdio.h>
#include <sys/mman.h>
#include <time.h>
#include <ucontext.h>

#include <set>
#include <vector>

namespace emu {

static std::set<const void *> host_allocations;
void * device_alloc(size_t bytes) { void * p = nullptr; return posix_memalign(&p, 256, bytes ? bytes : 256) == 0 ? p : nullptr; }
void   device_free(void * p) { free(p); }
void * host_alloc(size_t bytes) { void * p = device_alloc(bytes); if (p) host_allocations.insert(p); return p; }
void   host_free(void * p) { host_allocations.erase(p); free(p); }
bool   is_host_alloc(const void * p) { return host_allocations.count(p) != 0; }
double now_ms() { timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec * 1e3 + t.tv_nsec * 1e-6; }
unsigned long long global_timer_ns() { timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return (unsigned long long)t.tv_sec * 1000000000ull + (unsigned long long)t.tv_nsec; }

struct Warp {
	unsigned slots[32];
	unsigned arrived = 0;      // lanes that have deposited in the current rendezvous
	unsigned mask = 0;         // mask of the current rendezvous (all participants must name the same one)
	unsigned generation = 0;
	unsigned leaving = 0;      // lanes that still have to pick up their result of the finished rendezvous
};
struct Block { unsigned barrier_arrived = 0, barrier_generation = 0, alive = 0; };

struct Thread {
	ucontext_t context;
	void * stack = nullptr;
	bool done = false;
	uint3 thread_index, block_index;
	unsigned lane = 0;
	Warp * warp = nullptr;
	Block * block = nullptr;
};

Thread * current = nullptr;
static ucontext_t scheduler_context;
static dim3 g_block_dim, g_grid_dim;
static const std::function<void()> * g_body = nullptr;
static long long g_clock = 0;
static const size_t STACK_BYTES = 1 << 20;

uint3 & thread_idx() { return current->thread_index; }
uint3 & block_idx() { return current->block_index; }
dim3 & block_dim() { return g_block_dim; }
dim3 & grid_dim() { return g_grid_dim; }
long long clock() { return g_clock += 64; }

void yield() { g_clock++; swapcontext(&current->context, &scheduler_context); }

void sync_threads() {
	Block * b = current->block;
	const unsigned generation = b->barrier_generation;
	if (++b->barrier_arrived == b->alive) { b->barrier_arrived = 0; b->barrier_generation++; return; }
	while (b->barrier_generation == generation) yield();
}

unsigned warp_exchange(unsigned mask, unsigned value, int source_lane) {
	Warp * w = current->warp;
	const unsigned lane = current->lane;
	while (w->leaving != 0) yield();                  // the previous rendezvous is still being read
	if (w->arrived == 0) w->mask = mask;
	else if (w->mask != mask) { fprintf(stderr, "emu: lanes of one warp disagree on a collective's mask (%08x vs %08x)\n", w->mask, mask); abort(); }
	if (!((mask >> lane) & 1u)) { fprintf(stderr, "emu: lane %u calls a collective whose mask %08x does not name it\n", lane, mask); abort(); }
	w->slots[lane] = value;
	w->arrived |= 1u << lane;
	const unsigned generation = w->generation;
	if (w->arrived == mask) { w->leaving = mask; w->arrived = 0; w->generation++; }
	else while (w->generation == generation) yield();
	unsigned result = 0u;
	if (source_lane == -2) { for (int l = 0; l < 32; l++) if ((mask >> l) & 1u) result |= w->slots[l]; }   // ballot: OR of what the lanes deposited
	else if (source_lane >= 0) result = ((mask >> source_lane) & 1u) ? w->slots[source_lane] : value;
	w->leaving &= ~(1u << lane);
	return result;
}

static void fibre_main() {
	(*g_body)();
	current->done = true;
	current->block->alive--;
	Block * b = current->block;   // a thread that returns no longer takes part in __syncthreads
	if (b->alive != 0 && b->barrier_arrived == b->alive) { b->barrier_arrived = 0; b->barrier_generation++; }
	swapcontext(&current->context, &scheduler_context);
}

static void run(std::vector<Thread> & threads) {
	size_t remaining = threads.size();
	while (remaining) {
		for (Thread & t : threads) {
			if (t.done) continue;
			current = &t;
			swapcontext(&scheduler_context, &t.context);
			if (t.done) remaining--;
		}
	}
	current = nullptr;
}

void launch(dim3 grid, dim3 block, bool blocks_concurrent, const std::function<void()> & body) {
	g_block_dim = block; g_grid_dim = grid; g_body = &body;
	const unsigned per_block = block.x * block.y * block.z, blocks = grid.x * grid.y * grid.z;
	const unsigned warps_per_block = (per_block + 31) / 32;
	const unsigned batch = blocks_concurrent ? blocks : 1;

	std::vector<void *> stacks(size_t(batch) * per_block);
	for (void *& s : stacks) s = mmap(nullptr, STACK_BYTES, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);

	for (unsigned first = 0; first < blocks; first += batch) {
		const unsigned n = std::min(batch, blocks - first);
		std::vector<Thread> threads(size_t(n) * per_block);
		std::vector<Warp> warps(size_t(n) * warps_per_block);
		std::vector<Block> block_state(n);
		for (unsigned b = 0; b < n; b++) {
			block_state[b].alive = per_block;
			const unsigned id = first + b;
			for (unsigned t = 0; t < per_block; t++) {
				Thread & th = threads[size_t(b) * per_block + t];
				th.stack = stacks[size_t(b) * per_block + t];
				th.thread_index = { t % block.x, (t / block.x) % block.y, t / (block.x * block.y) };
				th.block_index = { id % grid.x, (id / grid.x) % grid.y, id / (grid.x * grid.y) };
				th.lane = t & 31;
				th.warp = &warps[size_t(b) * warps_per_block + t / 32];
				th.block = &block_state[b];
				getcontext(&th.context);
				th.context.uc_stack.ss_sp = th.stack;
				th.context.uc_stack.ss_size = STACK_BYTES;
				th.context.uc_link = nullptr;
				makecontext(&th.context, fibre_main, 0);
			}
		}
		run(threads);
	}
	for (void * s : stacks) munmap(s, STACK_BYTES);
	g_body = nullptr;
}

} // namespace emu
