from .interfaces import ExchangeTaskScheduler, ExchangeTaskSpec  # noqa: F401
from .pull_based import PullBasedShuffleTaskScheduler  # noqa: F401
from .push_based import PushBasedShuffleTaskScheduler  # noqa: F401
from .sort_task_spec import SortTaskSpec  # noqa: F401
from .shuffle_task_spec import ShuffleTaskSpec  # noqa: F401
from .aggregate_task_spec import SortAggregateTaskSpec  # noqa: F401
from .device_exchange import DeviceDataset, DeviceExchange  # noqa: F401
