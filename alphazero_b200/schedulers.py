"""Temperature schedules (alphazero/schedulers.py:4-46): step index -> temperature of the move selection."""
from .base import TemperatureScheduler


class ConstantTemperatureScheduler(TemperatureScheduler):
    """The reference's "constant" schedule returns temperature 0 at every step (schedulers.py:17-19)."""

    def __init__(self, temp_max_step: int, temp_min_step: int, max_steps: int) -> None:
        super().__init__(temp_max_step, temp_min_step, max_steps)
        if temp_min_step != temp_max_step:
            raise ValueError("temp_min_step should be equal to temp_max_step for constant scheduler.")

    def compute_temperature(self, step: int) -> float:
        return 0


class LinearTemperatureScheduler(TemperatureScheduler):
    """1 up to temp_max_step, 0 from temp_min_step on, linear in between (schedulers.py:33-40)."""

    def __init__(self, temp_max_step: int, temp_min_step: int, max_steps: int) -> None:
        super().__init__(temp_max_step, temp_min_step, max_steps)
        if temp_min_step < temp_max_step:
            raise ValueError("temp_min_step should be greater than temp_max_step for linear scheduler.")

    def compute_temperature(self, step: int) -> float:
        if step <= self.temp_max_step:
            return 1
        if step >= self.temp_min_step:
            return 0
        return 1 - (step - self.temp_max_step) / (self.temp_min_step - self.temp_max_step)


TEMP_SCHEDULERS = {"constant": ConstantTemperatureScheduler, "linear": LinearTemperatureScheduler}
