// cuda_worker_group.hpp -- the reference-side binding: an IWorkerGroup that hands herdsman's
// MapTask / ReduceTask to the B200 backend through the C ABI of include/herdb.h.
//
// This header is compiled INSIDE herdsman (C++20), next to GrpcWorkerGroup / LambdaWorkerGroup:
//   implements   IWorkerGroup                 include/execution/worker/i_worker_group.hpp:12-49
//   consumes     herd::common::MapTask        common/include/herd/common/model/worker/task.hpp:12-21
//                herd::common::Circuit        common/include/herd/common/model/circuit/circuit.hpp:14-30
//                DAG<node_t>                  common/include/herd/common/dag.hpp
//   selected in  build_worker_group()         src/main.cpp:67-84   (one more `else if`)
// It depends only on herdsman's own headers and on herdb.h / libherdb.so.  It is header-only so
// that herdsman's CMake needs one include directory and one link line.
//
// Completion contract: schedule_task returns a PENDING handle; a backend host thread completes
// it after the output partition file is in place.  The executor installs its callback only AFTER
// schedule_task has returned (src/execution/executor/executor.cpp:86-93) and
// TaskHandle::set_completion_callback is neither virtual nor re-firing (i_worker_group.cpp:4-20), so a
// completion that arrives in between would be dropped and the job would hang.  A task can fail within
// microseconds (missing key file, missing partition, refused circuit), so this is a real window, not a
// theoretical one.  The adapter closes it without touching herdsman:
//   * schedule_task is only ever called from the single executor thread (executor.cpp:46-80), and that thread
//     calls concurrent_workers() right after every schedule_task_on_worker (the loop condition of
//     schedule_tasks_on_workers, executor.cpp:96-113).  So whenever the executor thread ENTERS schedule_task
//     or concurrent_workers, the callback of every handle returned earlier is installed: those handles are
//     "armed".
//   * A completion of an armed handle fires at once on the backend thread.  A completion of a handle that is
//     not armed yet is parked; the executor thread's next entry (at most the time of one
//     set_completion_callback away) releases it, and a completer thread releases parked completions after a
//     bounded wait (kArmTimeout) for callers that never come back.
#ifndef HERD_B200_CUDA_WORKER_GROUP_HPP
#define HERD_B200_CUDA_WORKER_GROUP_HPP

#include <atomic>
#include <cassert>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <variant>
#include <vector>

#include "execution/worker/i_worker_group.hpp"
#include "herd/common/model/worker/task.hpp"
#include "herdb.h"

class CudaWorkerGroup final : public IWorkerGroup
{
public:
	class CudaTaskHandle final : public TaskHandle
	{
	public:
		[[nodiscard]] Status status() const noexcept override { return status_.load(); }
		void complete(Status status) { status_.store(status); mark_completed(); }
	private:
		friend class CudaWorkerGroup;
		std::atomic<Status> status_{Status::PENDING};
		uint64_t epoch_ = 0;                                             // value of group.next_epoch_ when the handle was returned
	};

	CudaWorkerGroup(const std::vector<int>& devices, const std::string& storage_directory,
	                const std::string& key_directory)
	{
		const int rc = herdb_create(&ctx_, devices.data(), static_cast<int>(devices.size()),
		                            storage_directory.c_str(), key_directory.c_str());
		if(rc != HERDB_OK)
		{
			throw std::runtime_error(std::string("CudaWorkerGroup: ") + herdb_strerror(rc));
		}
		concurrent_ = static_cast<std::size_t>(herdb_task_window(ctx_));    // task threads + a queue small tasks are coalesced from
		completer_ = std::thread([this]() { completer_body(); });
	}

	~CudaWorkerGroup() override
	{
		herdb_destroy(ctx_);                                             // waits for running tasks; their completions are parked or fired
		{
			std::unique_lock lock(arm_mutex_);
			shutting_down_ = true;
		}
		arm_cv_.notify_all();
		completer_.join();
	}

	CudaWorkerGroup(const CudaWorkerGroup&) = delete;
	CudaWorkerGroup& operator=(const CudaWorkerGroup&) = delete;

	std::size_t concurrent_workers() const noexcept override
	{
		const_cast<CudaWorkerGroup*>(this)->arm_returned_handles();      // executor thread: everything returned so far has its callback
		return concurrent_;
	}

	std::shared_ptr<TaskHandle> schedule_task(const herd::common::task_t& task) override
	{
		assert(!std::holds_alternative<std::monostate>(task));          // grpc_worker_group.cpp:74
		arm_returned_handles();
		auto handle = std::make_shared<CudaTaskHandle>();
		{
			std::unique_lock lock(arm_mutex_);
			handle->epoch_ = next_epoch_++;
		}
		auto* pending = new Pending{this, handle};
		{
			std::unique_lock lock(pending_mutex_);
			pending_[pending] = handle;                                  // keeps the handle alive (grpc_worker_group.cpp:100)
		}
		int rc = HERDB_OK;
		if(std::holds_alternative<herd::common::ReduceTask>(task))       // task.hpp:23-32
		{
			const auto& reduce_task = std::get<herd::common::ReduceTask>(task);
			Flat flat = flatten(reduce_task.circuit);
			std::vector<std::string> uuids;
			std::vector<herdb_input_frame> frames;
			uuids.reserve(reduce_task.input_data_frame_ptrs.size());
			for(const auto& ptr: reduce_task.input_data_frame_ptrs)
			{
				uuids.emplace_back(ptr.pointer.uuid.as_string());
				frames.push_back(herdb_input_frame{uuids.back().c_str(), ptr.pointer.partition, ptr.row_count});
			}
			rc = herdb_reduce_async(
					ctx_, reduce_task.session_uuid.as_string().c_str(), frames.data(), frames.size(),
					reduce_task.output_data_frame_ptr.uuid.as_string().c_str(), reduce_task.output_data_frame_ptr.partition,
					static_cast<int>(reduce_task.crypto_key_ptr.schema_type),          // key file = <int(SchemaType)>.key, key_service.cpp:28-31
					&flat.circuit, &CudaWorkerGroup::on_done, pending);
		}
		else
		{
			const auto& map_task = std::get<herd::common::MapTask>(task);
			Flat flat = flatten(map_task.circuit);
			rc = herdb_map_async(
					ctx_, map_task.session_uuid.as_string().c_str(),
					map_task.input_data_frame_ptr.pointer.uuid.as_string().c_str(), map_task.input_data_frame_ptr.pointer.partition,
					map_task.input_data_frame_ptr.row_count,
					map_task.output_data_frame_ptr.uuid.as_string().c_str(), map_task.output_data_frame_ptr.partition,
					static_cast<int>(map_task.crypto_key_ptr.schema_type),
					&flat.circuit, &CudaWorkerGroup::on_done, pending);
		}
		if(rc != HERDB_OK)
		{
			// refused synchronously (malformed circuit, ...): ERROR, delivered like any other completion
			on_done(pending, HERDB_TASK_ERROR);
		}
		return handle;
	}

private:
	struct Pending { CudaWorkerGroup* group; std::shared_ptr<CudaTaskHandle> handle; };

	struct Flat
	{
		std::vector<herdb_node> nodes;
		std::vector<herdb_edge> edges;
		std::vector<uint8_t> in_types, out_types;
		std::vector<uint32_t> in_offsets;
		herdb_circuit circuit{};
	};

	// herd::common::Circuit -> herdb_circuit.  Edges are emitted per child in parents() order so that
	// operand order survives (DAG::parents(), dag.hpp:208-230).
	static Flat flatten(const herd::common::Circuit& circuit)
	{
		using namespace herd::common;
		Flat flat;
		const auto& graph = circuit.circuit_graph;
		for(auto iter = std::begin(graph); iter != std::end(graph); ++iter)
		{
			const auto node = *iter;
			herdb_node out{};
			const node_t& value = node.value();
			if(std::holds_alternative<InputNode>(value))
			{
				const auto& n = std::get<InputNode>(value);
				out.kind = HERDB_NODE_INPUT; out.tuple = n.tuple_index; out.field = n.field_index; out.bit = n.bit_index;
			}
			else if(std::holds_alternative<OutputNode>(value))
			{
				const auto& n = std::get<OutputNode>(value);
				out.kind = HERDB_NODE_OUTPUT; out.field = n.field_index; out.bit = n.bit_index;
			}
			else if(std::holds_alternative<ConstantNode>(value))
			{
				out.kind = HERDB_NODE_CONSTANT; out.const_val = std::get<ConstantNode>(value).value ? 1 : 0;
			}
			else
			{
				out.kind = HERDB_NODE_OPERATION;
				out.op = static_cast<uint8_t>(std::get<OperationNode>(value).type);     // same order as herdb_operation
			}
			flat.nodes.push_back(out);
			for(const auto& parent: node.parents())
			{
				flat.edges.push_back(herdb_edge{static_cast<uint32_t>(parent.node_id()), static_cast<uint32_t>(node.node_id())});
			}
		}
		flat.in_offsets.push_back(0);
		for(const auto& tuple: circuit.inputs)
		{
			for(const auto type: tuple) flat.in_types.push_back(static_cast<uint8_t>(type));        // same order as herdb_data_type
			flat.in_offsets.push_back(static_cast<uint32_t>(flat.in_types.size()));
		}
		for(const auto& column: circuit.output) flat.out_types.push_back(static_cast<uint8_t>(column.data_type));
		flat.circuit.nodes = flat.nodes.data(); flat.circuit.n_nodes = flat.nodes.size();
		flat.circuit.edges = flat.edges.data(); flat.circuit.n_edges = flat.edges.size();
		flat.circuit.in_field_types = flat.in_types.data(); flat.circuit.in_tuple_offsets = flat.in_offsets.data();
		flat.circuit.n_tuples = circuit.inputs.size();
		flat.circuit.out_field_types = flat.out_types.data(); flat.circuit.n_out_fields = flat.out_types.size();
		return flat;
	}

	static constexpr std::chrono::milliseconds kArmTimeout{200};

	// called on the executor thread: handles returned before this point have their callback installed
	void arm_returned_handles()
	{
		std::deque<Parked> release;
		{
			std::unique_lock lock(arm_mutex_);
			armed_epoch_ = next_epoch_;
			release.swap(parked_);
		}
		for(auto& parked: release)
		{
			parked.handle->complete(parked.status);
		}
	}

	// backend thread (or schedule_task itself for a refused task): the task is finished
	static void on_done(void* user, int task_status)
	{
		auto* pending = static_cast<Pending*>(user);
		CudaWorkerGroup* group = pending->group;
		auto handle = pending->handle;
		{
			std::unique_lock lock(group->pending_mutex_);
			group->pending_.erase(pending);
		}
		delete pending;
		using Status = TaskHandle::Status;
		const Status status = task_status == HERDB_TASK_COMPLETED ? Status::COMPLETED
		                    : task_status == HERDB_TASK_TIME_OUT ? Status::TIME_OUT : Status::ERROR;
		{
			std::unique_lock lock(group->arm_mutex_);
			if(handle->epoch_ >= group->armed_epoch_)                      // callback possibly not installed yet: park
			{
				group->parked_.push_back(Parked{handle, status, std::chrono::steady_clock::now() + kArmTimeout});
				lock.unlock();
				group->arm_cv_.notify_all();
				return;
			}
		}
		handle->complete(status);
	}

	// releases parked completions whose bounded wait is over (a caller that never re-enters the group)
	void completer_body()
	{
		std::unique_lock lock(arm_mutex_);
		while(true)
		{
			if(parked_.empty())
			{
				if(shutting_down_) return;
				arm_cv_.wait(lock);
				continue;
			}
			const auto deadline = parked_.front().deadline;
			if(!shutting_down_ && std::chrono::steady_clock::now() < deadline)
			{
				arm_cv_.wait_until(lock, deadline);
				continue;
			}
			Parked parked = std::move(parked_.front());
			parked_.pop_front();
			lock.unlock();
			parked.handle->complete(parked.status);
			lock.lock();
		}
	}

	struct Parked
	{
		std::shared_ptr<CudaTaskHandle> handle;
		TaskHandle::Status status;
		std::chrono::steady_clock::time_point deadline;
	};

	herdb_ctx* ctx_ = nullptr;
	std::size_t concurrent_ = 1;
	std::mutex pending_mutex_;
	std::map<Pending*, std::shared_ptr<CudaTaskHandle>> pending_;
	std::mutex arm_mutex_;
	std::condition_variable arm_cv_;
	uint64_t next_epoch_ = 0, armed_epoch_ = 0;                          // handles with epoch < armed_epoch_ have their callback
	std::deque<Parked> parked_;
	bool shutting_down_ = false;
	std::thread completer_;
};

#endif //HERD_B200_CUDA_WORKER_GROUP_HPP
