// fake_executor.cpp -- drives include/herd_b200/cuda_worker_group.hpp the way herdsman's Executor does, LINKED
// against the reference's own IWorkerGroup::TaskHandle (src/execution/worker/i_worker_group.cpp), task / circuit /
// DAG model (common/include) and UUID (common/src/uuid.cpp), and against libherdb.so.  TEST INFRASTRUCTURE: built
// in the build container by tests/adapter/Makefile from the reference sources where they lie (output in
// tests/adapter/_ref/, git-ignored, travels to the GPU box like the other built artefacts).
//
// What it mirrors (src/execution/executor/executor.cpp:82-119):
//     schedule_task_on_worker:  handle = worker_group->schedule_task(task);  handle->set_completion_callback(...)
//     schedule_tasks_on_workers: for(; job_count < max(0, worker_group->concurrent_workers() - pending); ++job_count) ...
// i.e. the callback is installed AFTER schedule_task returned and concurrent_workers() is polled after every task.
//
// usage: fake_executor <storage_dir> <key_dir> <session> <in_df> <out_df> <partitions> <rows> <devices>
// schedules, in this order: one 8-bit-adder MapTask per partition; a MapTask on a partition that does not exist
// (fails within microseconds); a MapTask with a malformed circuit (refused inside schedule_task); a MapTask of a
// session without key file.  Prints one line per completion: "<label> <status>".  Exit code 0 iff every task
// reported exactly once and the statuses are COMPLETED... / ERROR / ERROR / ERROR.
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "herd_b200/cuda_worker_group.hpp"

using namespace herd::common;
using Status = IWorkerGroup::TaskHandle::Status;

// OR(AND(a, NOT b), AND(NOT a, b))
static DAG<node_t>::iterator xor_wire(DAG<node_t>& g, DAG<node_t>::iterator a, DAG<node_t>::iterator b)
{
	auto na = g.emplace(OperationNode{Operation::NOT}); g.add_edge(a, na);
	auto nb = g.emplace(OperationNode{Operation::NOT}); g.add_edge(b, nb);
	auto l = g.emplace(OperationNode{Operation::AND}); g.add_edge(a, l); g.add_edge(nb, l);
	auto r = g.emplace(OperationNode{Operation::AND}); g.add_edge(na, r); g.add_edge(b, r);
	auto o = g.emplace(OperationNode{Operation::OR}); g.add_edge(l, o); g.add_edge(r, o);
	return o;
}

// (UINT8 x, UINT8 y) -> UINT8 (x + y) mod 256, ripple carry in the AND/OR/NOT basis the wire can carry
static Circuit adder8()
{
	Circuit c;
	c.inputs = {{DataType::UINT8, DataType::UINT8}};
	c.output.emplace_back("sum", DataType::UINT8);
	auto& g = c.circuit_graph;
	std::vector<DAG<node_t>::iterator> x, y;
	for(unsigned b = 0; b < 8; ++b) x.push_back(g.emplace(InputNode{0, 0, b}));
	for(unsigned b = 0; b < 8; ++b) y.push_back(g.emplace(InputNode{0, 1, b}));
	bool have_carry = false;
	DAG<node_t>::iterator carry;
	for(unsigned b = 0; b < 8; ++b)
	{
		auto axb = xor_wire(g, x[b], y[b]);
		auto sum = have_carry ? xor_wire(g, axb, carry) : axb;
		if(b < 7)
		{
			auto ab = g.emplace(OperationNode{Operation::AND}); g.add_edge(x[b], ab); g.add_edge(y[b], ab);
			if(have_carry)
			{
				auto ac = g.emplace(OperationNode{Operation::AND}); g.add_edge(axb, ac); g.add_edge(carry, ac);
				auto co = g.emplace(OperationNode{Operation::OR}); g.add_edge(ab, co); g.add_edge(ac, co);
				carry = co;
			}
			else carry = ab;
			have_carry = true;
		}
		auto out = g.emplace(OutputNode{0, b}); g.add_edge(sum, out);
	}
	return c;
}

struct Events
{
	std::mutex mutex;
	std::condition_variable cv;
	std::map<std::string, std::vector<Status>> seen;
	void post(const std::string& label, Status status)
	{
		{ std::unique_lock lock(mutex); seen[label].push_back(status); }
		cv.notify_all();
	}
};

static const char* name(Status s)
{
	return s == Status::COMPLETED ? "COMPLETED" : s == Status::PENDING ? "PENDING" : s == Status::TIME_OUT ? "TIME_OUT" : "ERROR";
}

int main(int argc, char** argv)
{
	if(argc < 9) { std::fprintf(stderr, "usage: fake_executor storage key session in_df out_df partitions rows devices\n"); return 2; }
	const std::string storage = argv[1], key_dir = argv[2];
	const UUID session(argv[3]), in_df(argv[4]), out_df(argv[5]);
	const unsigned partitions = static_cast<unsigned>(std::atoi(argv[6]));
	const unsigned long rows = std::strtoul(argv[7], nullptr, 10);
	std::vector<int> devices;
	for(int d = 0; d < std::atoi(argv[8]); ++d) devices.push_back(d);

	std::shared_ptr<IWorkerGroup> worker_group = std::make_shared<CudaWorkerGroup>(devices, storage, key_dir);
	Events events;
	std::vector<std::shared_ptr<IWorkerGroup::TaskHandle>> handles;
	std::vector<std::pair<std::string, task_t>> tasks;
	const Circuit circuit = adder8();
	for(unsigned k = 0; k < partitions; ++k)
	{
		const unsigned long part_rows = rows / partitions + (k < rows % partitions ? 1 : 0);      // storage_service.cpp:121-129
		tasks.emplace_back("map" + std::to_string(k), MapTask{session, {{in_df, k}, part_rows}, {out_df, k}, {SchemaType::BINFHE}, circuit});
	}
	tasks.emplace_back("missing_partition", MapTask{session, {{in_df, 4242}, 1}, {out_df, 4242}, {SchemaType::BINFHE}, circuit});
	Circuit broken = circuit;
	broken.circuit_graph.emplace(OutputNode{7, 0});                                                  // an output field the circuit does not have
	tasks.emplace_back("malformed_circuit", MapTask{session, {{in_df, 0}, 1}, {out_df, 4243}, {SchemaType::BINFHE}, broken});
	tasks.emplace_back("no_key", MapTask{UUID("00000000-0000-4000-8000-00000000dead"), {{in_df, 0}, 1}, {out_df, 4244}, {SchemaType::BINFHE}, circuit});

	// Executor::schedule_tasks_on_workers + schedule_task_on_worker, with the window the group asks for
	std::size_t next = 0, finished = 0;
	while(finished < tasks.size())
	{
		std::size_t pending = next - finished;
		for(; next < tasks.size() && pending < worker_group->concurrent_workers(); ++pending, ++next)
		{
			const std::string label = tasks[next].first;
			auto task_handle = worker_group->schedule_task(tasks[next].second);
			task_handle->set_completion_callback([&events, label](Status status) { events.post(label, status); });
			handles.push_back(task_handle);
		}
		std::unique_lock lock(events.mutex);
		if(!events.cv.wait_for(lock, std::chrono::seconds(600), [&] {
			std::size_t n = 0; for(auto& kv: events.seen) n += kv.second.size(); return n > finished; }))
		{
			std::fprintf(stderr, "fake_executor: a completion was lost (waited 600 s)\n");
			return 3;
		}
		finished = 0; for(auto& kv: events.seen) finished += kv.second.size();
	}
	int rc = 0;
	for(auto& [label, task]: tasks)
	{
		const auto& st = events.seen[label];
		const Status want = label.rfind("map", 0) == 0 ? Status::COMPLETED : Status::ERROR;
		std::printf("%s %s\n", label.c_str(), st.size() == 1 ? name(st[0]) : "LOST_OR_DUPLICATED");
		if(st.size() != 1 || st[0] != want) rc = 1;
	}
	for(auto& h: handles) if(!h->completed()) rc = 1;
	return rc;
}
