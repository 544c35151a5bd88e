"""The reduce trees herdsman builds for a ReduceStage, restated from
ExecutionService::ReduceStageProgress (src/service/execution_service.cpp:560-689) so that tests can replay the
exact ReduceTask sequence -- including the labels the reference puts on its intermediate frames -- against the
backend.  TEST INFRASTRUCTURE.

A tree node is (data_frame, row_count_label, partition).  A task = (input nodes, output node), issued when all
its inputs exist (mark_task_completed, execution_service.cpp:691-705)."""


def build(policy, partitions, part_rows, in_df, out_df, hidden_df, per_node_count=2):
    """returns the list of tasks in a valid execution order: [(inputs, output)], inputs = [(df, row_count, part)]"""
    inputs = [(in_df, part_rows[i], i) for i in range(partitions)]
    output = (out_df, 1, 1)                      # `(output_data_frame, 1, 1)`: row_count 1, PARTITION 1 (execution_service.cpp:596,604,673)
    tasks = []
    if policy == "SEQUENCED":                    # :594-603 one task over every input partition
        tasks.append((inputs, output))
    elif policy == "PARALLEL":                   # :604-627 one reduce per partition into a hidden frame, then one over the hidden frame
        hidden = [(hidden_df, partitions, i) for i in range(partitions)]       # label row_count = partitions; each holds ONE row
        for i in range(partitions):
            tasks.append(([inputs[i]], hidden[i]))
        tasks.append((hidden, output))
    elif policy == "PARALLEL_FULL":              # :628-688 per-partition reduce, then groups of per_node_count level by level
        idx = 0
        layer = []
        for i in range(partitions):
            node = (hidden_df, 1, idx); idx += 1
            tasks.append(([inputs[i]], node))
            layer.append(node)
        while len(layer) > per_node_count:
            prev, layer = layer, []
            full = len(prev) - len(prev) % per_node_count
            for i in range(0, full, per_node_count):
                node = (hidden_df, 1, idx); idx += 1
                tasks.append((prev[i:i + per_node_count], node))
                layer.append(node)
            layer.extend(prev[full:])
        tasks.append((layer, output))
    else:
        raise ValueError(policy)
    return tasks


def hidden_partitions(policy, partitions, per_node_count=2):
    """partition count the reference gives the hidden frame (node_sum, execution_service.cpp:630-650)"""
    if policy == "PARALLEL":
        return partitions
    if policy == "PARALLEL_FULL":
        cur = total = partitions
        while cur > per_node_count:
            cur = cur // per_node_count + cur % per_node_count
            total += cur
        return total
    return 0
