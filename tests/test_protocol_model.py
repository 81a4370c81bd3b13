"""The flag/epoch protocols of the fused kernels, checked over EVERY interleaving of the ranks
(gpu_mpi_tests_b200/utils/protocol_model.py).  The negative cases matter as much as the positive
ones: they show the checker can see the race each design element exists to prevent."""
import pytest

from gpu_mpi_tests_b200.utils.protocol_model import (ProtocolViolation, allgather_model, check, halo_model,
                                                      ll_allreduce_model)


@pytest.mark.parametrize("n_ranks,steps", [(2, 4), (3, 3)])
def test_halo_exchange_with_parity_double_buffer_is_safe(n_ranks, steps):
    states = check(halo_model, n_ranks=n_ranks, steps=steps, buffers=2)
    assert states > 50   # it really explored something


def test_halo_exchange_single_buffer_races():
    """One inbox per side is not enough: a neighbour that is a launch ahead overwrites ghosts."""
    with pytest.raises(ProtocolViolation, match="ghosts hold") as info:
        check(halo_model, n_ranks=2, steps=2, buffers=1)
    # the counterexample has one rank pushing step 2 before the other has read step 1
    names = [a for _r, a in info.value.trace]
    assert any(a.startswith("push2") for a in names) and names[-1] == "read1"


@pytest.mark.parametrize("n_ranks,calls", [(2, 3), (3, 2)])
def test_inplace_allgather_two_rounds_is_safe(n_ranks, calls):
    assert check(allgather_model, n_ranks=n_ranks, calls=calls, entry_round=True) > 100


def test_inplace_allgather_without_entry_round_races():
    """Storing into a peer's buffer before the peer has entered the call clobbers data the peer's
    user code is still reading from the previous call."""
    with pytest.raises(ProtocolViolation, match="holds") as info:
        check(allgather_model, n_ranks=2, calls=2, entry_round=False)
    assert any(a.startswith("store2") for _r, a in info.value.trace)


@pytest.mark.parametrize("n_ranks,calls", [(2, 4), (3, 3)])
def test_ll_allreduce_with_two_cell_sets_is_safe(n_ranks, calls):
    assert check(ll_allreduce_model, n_ranks=n_ranks, calls=calls, buffers=2) > 50


def test_ll_allreduce_single_cell_set_loses_values():
    """With one cell set a rank that is a call ahead replaces a value its peer has not taken yet;
    cells are matched by equality with the epoch, so the peer's wait can never end."""
    with pytest.raises(ProtocolViolation, match="deadlock"):
        check(ll_allreduce_model, n_ranks=2, calls=2, buffers=1)


def test_gather_between_reductions_must_not_let_senders_run_ahead():
    """Why gmt_gather_f64 ends with an all-to-all round although only the root needs the data: the
    reductions' cells are double-buffered by epoch parity, which holds only if no collective lets a
    rank get more than one call ahead of another."""
    with pytest.raises(ProtocolViolation, match="deadlock"):
        check(ll_allreduce_model, n_ranks=3, calls=3, buffers=2, gather_between="nowait")
    assert check(ll_allreduce_model, n_ranks=3, calls=3, buffers=2, gather_between="barrier") > 50


@pytest.mark.parametrize("n_ranks,calls", [(2, 4), (3, 2)])
def test_ll_allgather_in_place_needs_no_entry_round(n_ranks, calls):
    """Small in-place allgathers through LL cells: nobody writes another rank's user buffer, so the user
    always reads the current call's values although there is no "entered" round; one cell set is not
    enough (same lapping as the all-reduce)."""
    from gpu_mpi_tests_b200.utils.protocol_model import ll_allgather_model
    assert check(ll_allgather_model, n_ranks=n_ranks, calls=calls, buffers=2) > 50
    with pytest.raises(ProtocolViolation):
        check(ll_allgather_model, n_ranks=2, calls=3, buffers=1)


def test_pipelined_host_feed_needs_two_inboxes():
    """step_e2e_pipelined(): the host is one step ahead of the kernels.  With the host inbox
    double-buffered by epoch parity every kernel reads its own step's boundary values; with a single
    inbox the upload of the next step can land between two reads of the running kernel.  The
    synchronous call (no look-ahead) is safe with one inbox."""
    from gpu_mpi_tests_b200.utils.protocol_model import host_feed_model
    assert check(host_feed_model, steps=5, buffers=2, lookahead=1) > 20
    assert check(host_feed_model, steps=4, buffers=1, lookahead=0) > 10
    with pytest.raises(ProtocolViolation, match="sees the boundary values of step"):
        check(host_feed_model, steps=3, buffers=1, lookahead=1)
    # two steps of look-ahead would break parity double-buffering as well
    with pytest.raises(ProtocolViolation, match="sees the boundary values of step"):
        check(host_feed_model, steps=4, buffers=2, lookahead=2)


def test_checker_reports_deadlock():
    from gpu_mpi_tests_b200.utils.protocol_model import explore
    never = ("wait", lambda mem, p: False)
    with pytest.raises(ProtocolViolation, match="deadlock"):
        explore(1, (0,), [[never]])
