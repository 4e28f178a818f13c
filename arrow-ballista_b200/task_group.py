"""Host mirror of the task-group optimizer rule, for the plan nodes of the hot path
(/root/reference/ballista/core/src/physical_optimizer/task_group.rs:35-168). It is what puts the `CoalesceTasksExec` of
SURVEY.md §8 a7 where `GpuExecutionEngine` expects it: the scheduler runs it over a stage plan before a task group is
shipped (/root/reference/ballista/scheduler/src/state/task_manager.rs:223-226).

  transform_down (:44-75)   the stage's ShuffleWriterExec takes the task group's partitions (`with_partitions`); a
                            CoalesceTasksExec is put above every leaf and above every Final / FinalPartitioned aggregate
                            (:143-151; the reference also lists Union, SortPreservingMerge, CoalescePartitions and HashJoin,
                            which have no mirror here), and the walk does not descend below an inserted node;
  transform_up (:77-110)    a "mapping" node (:154-160: Filter, CoalesceBatches, Projection, LocalLimit, Partial aggregate -
                            here CoalesceBatchesExec and the Partial aggregate) whose only child is a CoalesceTasksExec is
                            pushed below it, so that the fan-in ends up directly under the writer.
"""
from .aggregate import AggregateMode
from .codec import GpuAggregateExecNode
from .shuffle_reader import CoalesceBatchesExec, CoalesceTasksExec
from .shuffle_writer import GpuShuffleWriterExec


def _is_aggregation(node, *modes):
    return isinstance(node, GpuAggregateExecNode) and node.mode in modes


def insert_coalesce(node):
    """task_group.rs:143-151, for the nodes that exist in this mirror."""
    return not node.children() or _is_aggregation(node, AggregateMode.Final, AggregateMode.FinalPartitioned)


def is_mapping(node):
    """task_group.rs:154-160, for the nodes that exist in this mirror."""
    return isinstance(node, CoalesceBatchesExec) or _is_aggregation(node, AggregateMode.Partial)


class OptimizeTaskGroup:
    def __init__(self, partitions):
        self.partitions = [int(p) for p in partitions]

    def name(self):
        return "optimize_task_group"

    def optimize(self, plan):
        return self._transform_up(self._transform_down(plan))

    def _transform_down(self, node):
        if isinstance(node, GpuShuffleWriterExec):
            node = node.with_partitions(self.partitions)
        elif insert_coalesce(node):
            return CoalesceTasksExec(node, self.partitions)
        children = node.children()
        new = [self._transform_down(c) for c in children]
        return node.with_new_children(new) if any(a is not b for a, b in zip(new, children)) else node

    def _transform_up(self, node):
        children = node.children()
        if children:
            new = [self._transform_up(c) for c in children]
            if any(a is not b for a, b in zip(new, children)):  # with_new_children_if_necessary
                node = node.with_new_children(new)
        children = node.children()
        if is_mapping(node) and len(children) == 1 and isinstance(children[0], CoalesceTasksExec):
            return CoalesceTasksExec(node.with_new_children(children[0].children()), self.partitions)
        return node
