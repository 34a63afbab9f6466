from .combinable import (Combinable, CombinationLogic, CountLogic, TimerLogic,
                         all_should_send, any_should_send)

__all__ = ["Combinable", "CombinationLogic", "CountLogic", "TimerLogic",
           "all_should_send", "any_should_send"]
