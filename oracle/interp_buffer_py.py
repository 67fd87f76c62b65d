"""TEST INFRASTRUCTURE — literal restatement of the reference's ba::InterpolationBufferT
(reference include/ba/InterpolationBuffer.h:38-210), the buffer whose GetRange feeds AddImuResidual.

PARITY UNPINNED like the rest of oracle/: the reference ships no test or fixture for this class; the
restatement follows the source statement by statement (guess from average_dt, then the element-by-element
walk), which is what include/ba/InterpolationBuffer.h (this repo; binary search) is compared with.
Elements are rows [w0 w1 w2 a0 a1 a2 time]; ImuMeasurementT's operator* / operator+ scale and add w and a
(reference include/ba/Types.h:222-244).  Only tests/ may import this module.
"""
import numpy as np


class InterpolationBufferRef:
    def __init__(self):
        self.elements = []
        self.start_time = self.end_time = self.average_dt = -1.0

    # InterpolationBuffer.h:64-77
    def AddElement(self, row):
        row = np.asarray(row, dtype=np.float64)
        assert row[6] > self.end_time
        n = len(self.elements)
        dt = 0.0 if n == 0 else row[6] - self.elements[-1][6]
        self.average_dt = dt if self.average_dt == -1 else (self.average_dt * n + dt) / (n + 1)
        self.elements.append(row)
        self.end_time = row[6]
        self.start_time = self.elements[0][6]

    # InterpolationBuffer.h:111-113
    def HasElement(self, time):
        return self.start_time <= time <= self.end_time

    @staticmethod
    def _lerp(a, b, s, time):
        res = a * (1 - s) + b * s
        res[6] = time
        return res

    # InterpolationBuffer.h:134-182 (queries outside [start_time, end_time) are not restated: the source reads
    # out of bounds there)
    def GetElement(self, time):
        e = self.elements
        n = len(e)
        guess = int((time - self.start_time) / self.average_dt) if self.average_dt > 0 else 0
        guess = min(max(guess, 0), n - 1)
        if e[guess][6] > time:
            if guess == 0:
                return e[0].copy(), 0
            while (guess - 1) > 0 and e[guess - 1][6] > time:
                guess -= 1
            s = (time - e[guess - 1][6]) / (e[guess][6] - e[guess - 1][6])
            return self._lerp(e[guess - 1], e[guess], s, time), guess - 1
        if guess == n - 1:
            return e[-1].copy(), guess
        while (guess + 1) < n and e[guess + 1][6] < time:
            guess += 1
        s = (time - e[guess][6]) / (e[guess + 1][6] - e[guess][6])
        return self._lerp(e[guess], e[guess + 1], s, time), guess

    # InterpolationBuffer.h:89-104 -> (returned flag, index, element)
    def GetNext(self, max_time, index):
        e = self.elements
        if index + 1 >= len(e) or e[index + 1][6] > max_time:
            out, index = self.GetElement(max_time)
            return False, index, out
        index += 1
        return True, index, e[index].copy()

    # InterpolationBuffer.h:184-210
    def GetRange(self, start, end):
        out = []
        if start < self.start_time:
            start = self.start_time
        if end > self.end_time:
            end = self.end_time
        if self.HasElement(start):
            first, index = self.GetElement(start)
            out.append(first)
            while True:
                more, index, meas = self.GetNext(end, index)
                if not more:
                    break
                out.append(meas)
            out.append(meas)
        return np.array(out).reshape(-1, 7)
