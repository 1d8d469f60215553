{{/* name helper: override, else chart name; 24 chars max */}}
{{- define "fx.name" -}}
{{- default .Chart.Name .Values.nameOverride | trunc 24 | trimSuffix "-" -}}
{{- end -}}

{{- define "fx.labels" -}}
app.kubernetes.io/name: {{ include "fx.name" . }}
app.kubernetes.io/version: {{ .Chart.AppVersion | quote }}
{{- range $k, $v := .Values.extraLabels }}
{{ $k }}: {{ $v | quote }}
{{- end }}
{{- end -}}

{{- define "fx.userdata" -}}
#cloud-config
hostname: {{ include "fx.name" . }}
ssh_authorized_keys:
  - {{ .Values.sshKey }}
write_files:
  - path: /etc/motd
    content: |
{{ .Values.motd | indent 6 }}
runcmd:
  - [ sh, -c, "echo {{ .Release.Name }} > /etc/release" ]
{{- end -}}
