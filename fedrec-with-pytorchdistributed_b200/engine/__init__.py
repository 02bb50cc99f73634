from .step import TrainStep  # noqa: F401
from .trainer import Trainer, train_on_step, process_news_grad, process_user_grad, validate_impressions  # noqa: F401
from .federated import FederatedRunner, run_entrypoint  # noqa: F401
from .serving import Recommender  # noqa: F401
