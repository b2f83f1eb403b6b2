"""Drop-in for the reference's utils/camera.py (ImageFeatureEncoder, utils/camera.py:6-65).

Only the boundary is in scope: the attribute contract the fusion model reads (`patch_size`,
`resize_height`, `resize_width`) and the token layout (B, M, 384), row-major y*grid_w+x.  The DINO
backbone stays the stock Hugging Face module.  Two differences, both on the hand-off:
tokens stay on the device instead of a per-image .cpu() (set keep_on_device=False for the
reference behaviour) and can be handed over in 16 bits (token_dtype), and the dinov2 branch defines resize_height/resize_width (224), which the
reference's forward needs but its constructor forgets.
"""
import torch


class ImageFeatureEncoder:
    def __init__(self, config, device="cuda", keep_on_device=True, token_dtype=None):
        self.device = device
        self.keep_on_device = keep_on_device
        # token_dtype=torch.bfloat16 / torch.float16 hands the tokens over in 16 bits: half the bytes every
        # point gathers; the fusion kernels accept fp32, fp16 and bf16 tokens.  None keeps what the backbone returns.
        self.token_dtype = token_dtype
        self.model_name = config['train_params']['image_encoder'].lower()
        print("===========================================================")
        print(f"[INFO] Initializing ImageFeatureEncoder: {self.model_name}")
        print(f"[INFO] Using device: {self.device}")
        print("===========================================================")
        from transformers import AutoModel, AutoImageProcessor
        if self.model_name == "dinov2":
            model_id = "facebook/dinov2-small"
            self.processor = AutoImageProcessor.from_pretrained(model_id, use_fast=True)
            self.patch_size = 14
            self.resize_size = 224
            self.resize_height = self.resize_width = 224
            self.n_register = 0
        elif self.model_name == "dinov3":
            model_id = "facebook/dinov3-vits16-pretrain-lvd1689m"
            self.processor = AutoImageProcessor.from_pretrained(model_id, use_fast=True)
            self.patch_size = 16
            self.processor.size = {"height": 640, "width": 1138}
            self.processor.do_center_crop = False
            self.processor.do_resize = True
            self.resize_height = self.processor.size["height"]
            self.resize_width = self.processor.size["width"]
            self.n_register = 4
        else:
            raise ValueError(f"Unsupported model name: {self.model_name}")
        self.model = AutoModel.from_pretrained(model_id).to(self.device)
        self.model.eval()

    def __call__(self, images):
        dev_type = self.device.type if isinstance(self.device, torch.device) else torch.device(self.device).type
        patch, cls = [], []
        with torch.inference_mode(), torch.autocast(device_type=dev_type, dtype=torch.float16):
            for img in images:
                inputs = self.processor(images=img, return_tensors="pt").to(self.device)
                feats = self.model(**inputs).last_hidden_state.squeeze(0)
                tokens = feats[1:-4]  # the reference drops CLS and the last four (register) tokens
                if self.token_dtype is not None:
                    tokens = tokens.to(self.token_dtype)
                patch.append(tokens if self.keep_on_device else tokens.cpu())
                cls.append(feats[0] if self.keep_on_device else feats[0].cpu())
        return {"patch_features": torch.stack(patch), "global_features": torch.stack(cls)}
