"""Command-line entry point with the reference's interface:

    python main.py --batch_size 128 --epochs 20 --world_size 8 --save_path ./run/

Same flags as ref:main.py:21-28 (plus the additions in pmdist/cli.py); spawns one process per GPU
(ref:main.py:185-188) and writes train.log / test.log / model_{epoch}.pth / the two PNG curves.
"""
from pmdist.train import main

if __name__ == '__main__':
    main()
