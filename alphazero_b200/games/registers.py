"""Name -> class maps with the reference's keys (alphazero/games/registers.py:11-50); Othello is the game the B200
engine implements (Connect4 / TicTacToe stay on the reference's Python path)."""
from ..base import DataTransf, MoveFormat, dotdict
from .othello import OthelloBoard, OthelloConfig, OthelloNet

GAMES_SET = {"othello"}
CONFIGS_REGISTER = {"othello": OthelloConfig}
BOARDS_REGISTER = {"othello": OthelloBoard}
NETWORKS_REGISTER = {"othello": OthelloNet}
MOVE_FORMATS_REGISTER = {"othello": MoveFormat.ROW_COL}
DATA_AUGMENT_STRATEGIES = {
    "othello": dotdict({"reflection": DataTransf.REFLECT_H,
                        "rotations": [DataTransf.ROTATE_90, DataTransf.ROTATE_180, DataTransf.ROTATE_270]}),
}
