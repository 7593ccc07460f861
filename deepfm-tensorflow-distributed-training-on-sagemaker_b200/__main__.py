"""``python -m deepfm_b200 --task_type train ...`` (reference: `tf.app.run()` HVD:433-435, PS:469-471)."""
from deepfm_b200.runtime.cli import main

if __name__ == "__main__":
    raise SystemExit(main())
