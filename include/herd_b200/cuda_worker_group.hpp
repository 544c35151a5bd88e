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
// Completion contract: schedule_task returns a PENDING handle; the backend's host
// thread fires mark_completed() after the output partition file is in place.  The executor
// installs its callback after schedule_task returns (src/execution/executor/executor.cpp:86-93);
// TaskHandle::set_completion_callback is not virtual, so -- exactly like GrpcWorkerGroup -- a task
// that completes before the callback is installed would be lost.  The backend therefore never
// completes a task from inside schedule_task, and completion of real tasks takes milliseconds at
// least; for the malformed-task case (refused synchronously) completion is deferred to a detached
// thread that waits for the callback to be installed.
#ifndef HERD_B200_CUDA_WORKER_GROUP_HPP
#define HERD_B200_CUDA_WORKER_GROUP_HPP

#include <atomic>
#include <cassert>
#include <chrono>
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
		std::atomic<Status> status_{Status::PENDING};
	};

	CudaWorkerGroup(const std::vector<int>& devices, const std::string& storage_directory,
	                const std::string& key_directory, std::size_t concurrent_tasks = 2)
		: concurrent_(concurrent_tasks)
	{
		const int rc = herdb_create(&ctx_, devices.data(), static_cast<int>(devices.size()),
		                            storage_directory.c_str(), key_directory.c_str());
		if(rc != HERDB_OK)
		{
			throw std::runtime_error(std::string("CudaWorkerGroup: ") + herdb_strerror(rc));
		}
	}

	~CudaWorkerGroup() override { herdb_destroy(ctx_); }

	CudaWorkerGroup(const CudaWorkerGroup&) = delete;
	CudaWorkerGroup& operator=(const CudaWorkerGroup&) = delete;

	std::size_t concurrent_workers() const noexcept override { return concurrent_; }

	std::shared_ptr<TaskHandle> schedule_task(const herd::common::task_t& task) override
	{
		assert(!std::holds_alternative<std::monostate>(task));          // grpc_worker_group.cpp:74
		auto handle = std::make_shared<CudaTaskHandle>();
		auto* pending = new Pending{this, handle};
		{
			std::unique_lock lock(pending_mutex_);
			pending_[pending] = handle;                                  // keeps the handle alive (grpc_worker_group.cpp:100)
		}
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
			const int rc = herdb_reduce_async(
					ctx_, reduce_task.session_uuid.as_string().c_str(), frames.data(), frames.size(),
					reduce_task.output_data_frame_ptr.uuid.as_string().c_str(), reduce_task.output_data_frame_ptr.partition,
					&flat.circuit, &CudaWorkerGroup::on_done, pending);
			if(rc != HERDB_OK)
			{
				refuse(pending, handle);
			}
			return handle;
		}
		const auto& map_task = std::get<herd::common::MapTask>(task);
		Flat flat = flatten(map_task.circuit);
		const int rc = herdb_map_async(
				ctx_, map_task.session_uuid.as_string().c_str(),
				map_task.input_data_frame_ptr.pointer.uuid.as_string().c_str(), map_task.input_data_frame_ptr.pointer.partition,
				map_task.input_data_frame_ptr.row_count,
				map_task.output_data_frame_ptr.uuid.as_string().c_str(), map_task.output_data_frame_ptr.partition,
				&flat.circuit, &CudaWorkerGroup::on_done, pending);
		if(rc != HERDB_OK)
		{
			refuse(pending, handle);
		}
		return handle;
	}

private:
	struct Pending { CudaWorkerGroup* group; std::shared_ptr<CudaTaskHandle> handle; };

	// the backend refused the task synchronously (malformed circuit, ...): ERROR, but not from inside schedule_task
	void refuse(Pending* pending, const std::shared_ptr<CudaTaskHandle>& handle)
	{
		{
			std::unique_lock lock(pending_mutex_);
			pending_.erase(pending);
		}
		delete pending;
		fail_later(handle);
	}
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

	static void on_done(void* user, int task_status)
	{
		auto* pending = static_cast<Pending*>(user);
		{
			std::unique_lock lock(pending->group->pending_mutex_);
			pending->group->pending_.erase(pending);
		}
		using Status = TaskHandle::Status;
		pending->handle->complete(task_status == HERDB_TASK_COMPLETED ? Status::COMPLETED
		                        : task_status == HERDB_TASK_TIME_OUT ? Status::TIME_OUT : Status::ERROR);
		delete pending;
	}

	// a task refused synchronously must still complete asynchronously (see the header comment)
	static void fail_later(std::shared_ptr<CudaTaskHandle> handle)
	{
		std::thread([handle = std::move(handle)]()
		{
			std::this_thread::sleep_for(std::chrono::milliseconds(50));
			handle->complete(TaskHandle::Status::ERROR);
		}).detach();
	}

	herdb_ctx* ctx_ = nullptr;
	std::size_t concurrent_;
	std::mutex pending_mutex_;
	std::map<Pending*, std::shared_ptr<CudaTaskHandle>> pending_;
};

#endif //HERD_B200_CUDA_WORKER_GROUP_HPP
