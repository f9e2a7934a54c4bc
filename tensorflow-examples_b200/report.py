"""The exact lines the reference scripts print."""


def simple_report(W, b, loss):
    """simple/simple.py:39 -- numpy str() of a [1] float32 array, twice, and of a float32 scalar."""
    return "W: %s b: %s loss %s" % (W, b, loss)


def progress_line(step, epoch, batch, batch_count, cost, elapsed_s, frequency):
    """distributed.py:157-161: five print() arguments joined by single spaces (hence the double
    spaces); the ms figure always divides by `frequency`, as the reference does."""
    return " ".join(["Step so far: %d," % (step + 1),
                     " Epoch so far: %2d," % (epoch + 1),
                     " Batch used: %3d of %3d," % (batch + 1, batch_count),
                     " Cost now: %.4f," % cost,
                     " Time spent (delta): %3.1fms" % float(elapsed_s * 1000 / frequency)])


def accuracy_line(acc):
    return "Acc: %2.2f" % acc                                   # distributed.py:164


def time_taken_line(seconds, strict_reference_bugs=False):
    """distributed.py:165 is `"Time Taken: 2fs" % float(...)`, which raises TypeError as shipped."""
    if strict_reference_bugs:
        return "Time Taken: 2fs" % float(seconds)               # raises, exactly like the reference
    return "Time Taken: %2fs" % float(seconds)


def final_cost_line(cost):
    return "Final Computed Cost: %.2f" % cost                   # distributed.py:166
